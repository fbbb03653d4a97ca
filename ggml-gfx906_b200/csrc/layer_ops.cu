// layer_ops.cu — the f32 ops that sit between the quantized MUL_MATs of a decoder layer, so that a Llama graph stays on
// the device instead of bouncing to the CPU backend between our mat-muls: RMS_NORM, ADD / MUL with ggml
// broadcasting, GLU (SWIGLU, fused or split), ROPE (normal and NeoX, YaRN, frequency factors), SOFT_MAX (scale, mask,
// ALiBi, sinks).  Definitions of truth: the CPU operators of the reference (src/ggml-cpu/ops.cpp:4138-4186 rms_norm,
// :3551-3610 swiglu, :6049-6330 rope, :5685-5790 soft_max; binary broadcast: src/ggml-cpu/binary-ops.cpp); the reference's
// CUDA counterparts are norm.cu, binbcast.cu, unary.cu, rope.cu, softmax.cu.  All are HBM-bound elementwise / row
// kernels on f32; tensor conventions are ggml's (ne in elements, nb in bytes, dim 0 contiguous).
#include "common.cuh"

namespace b200 {

struct t4 { int64_t ne[4]; int64_t nb[4]; };

__device__ __forceinline__ float block_sum(float v, float * red) {
    v = warp_sum(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    if (lane == 0) red[warp] = v;
    __syncthreads();
    v = lane < nw ? red[lane] : 0.0f;
    v = warp_sum(v);
    __syncthreads();
    return v;
}
__device__ __forceinline__ float block_max(float v, float * red) {
    v = warp_max(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    if (lane == 0) red[warp] = v;
    __syncthreads();
    v = lane < nw ? red[lane] : -INFINITY;
    v = warp_max(v);
    __syncthreads();
    return v;
}

// ---------------------------------------------------------------- RMS_NORM: y = x / sqrt(mean(x^2) + eps), one CTA per row
__global__ void __launch_bounds__(256) k_rms_norm(const uint8_t * __restrict__ x, const t4 s, float * __restrict__ y, float eps) {
    __shared__ float red[32];
    const int64_t row = blockIdx.x;
    const int64_t i1 = row % s.ne[1], i2 = (row / s.ne[1]) % s.ne[2], i3 = row / (s.ne[1] * s.ne[2]);
    const float * xr = reinterpret_cast<const float *>(x + i1 * s.nb[1] + i2 * s.nb[2] + i3 * s.nb[3]);
    float * yr = y + row * s.ne[0];
    float sum = 0.0f;
    for (int64_t i = threadIdx.x; i < s.ne[0]; i += blockDim.x) { const float v = xr[i]; sum += v * v; }
    sum = block_sum(sum, red);
    const float scale = 1.0f / sqrtf(sum / (float) s.ne[0] + eps);
    for (int64_t i = threadIdx.x; i < s.ne[0]; i += blockDim.x) yr[i] = xr[i] * scale;
}

// ---------------------------------------------------------------- ADD / MUL with broadcasting of b over a (dst has a's shape, contiguous)
template <int OP>
__global__ void __launch_bounds__(256) k_bin_bcast(const uint8_t * __restrict__ a, const t4 sa, const uint8_t * __restrict__ b, const t4 sb,
                                                   float * __restrict__ dst, int64_t n) {
    const int64_t i = (int64_t) blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const int64_t i0 = i % sa.ne[0], r = i / sa.ne[0];
    const int64_t i1 = r % sa.ne[1], i2 = (r / sa.ne[1]) % sa.ne[2], i3 = r / (sa.ne[1] * sa.ne[2]);
    const float va = *reinterpret_cast<const float *>(a + i0 * sa.nb[0] + i1 * sa.nb[1] + i2 * sa.nb[2] + i3 * sa.nb[3]);
    const float vb = *reinterpret_cast<const float *>(b + (i0 % sb.ne[0]) * sb.nb[0] + (i1 % sb.ne[1]) * sb.nb[1] +
                                                          (i2 % sb.ne[2]) * sb.nb[2] + (i3 % sb.ne[3]) * sb.nb[3]);
    dst[i] = OP == 0 ? va + vb : va * vb;
}

// ---------------------------------------------------------------- SWIGLU: y = silu(x) * g, rows of nc values
__global__ void __launch_bounds__(256) k_swiglu(const uint8_t * __restrict__ x, int64_t x_row_stride, const uint8_t * __restrict__ g, int64_t g_row_stride,
                                                float * __restrict__ y, int64_t nc, int64_t n) {
    const int64_t i = (int64_t) blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const int64_t row = i / nc, c = i - row * nc;
    const float xv = reinterpret_cast<const float *>(x + row * x_row_stride)[c];
    const float gv = reinterpret_cast<const float *>(g + row * g_row_stride)[c];
    y[i] = (xv / (1.0f + expf(-xv))) * gv;
}

// ---------------------------------------------------------------- ROPE (normal / NeoX), one thread per rotated pair
struct rope_args {
    t4 s;                      // src
    int64_t dnb[4];            // dst strides (bytes)
    int n_dims, neox;
    float freq_scale, ext_factor, attn_factor, theta_scale;
    float corr0, corr1;
    const int32_t * pos;
    const float * freq_factors;
};
__global__ void __launch_bounds__(256) k_rope(const uint8_t * __restrict__ x, uint8_t * __restrict__ y, const rope_args a) {
    const int64_t half = a.s.ne[0] / 2;
    const int64_t i = (int64_t) blockIdx.x * 256 + threadIdx.x;
    const int64_t total = half * a.s.ne[1] * a.s.ne[2] * a.s.ne[3];
    if (i >= total) return;
    const int64_t p = i % half, r = i / half;
    const int64_t i1 = r % a.s.ne[1], i2 = (r / a.s.ne[1]) % a.s.ne[2], i3 = r / (a.s.ne[1] * a.s.ne[2]);
    const uint8_t * xs = x + i1 * a.s.nb[1] + i2 * a.s.nb[2] + i3 * a.s.nb[3];
    uint8_t * yd = y + i1 * a.dnb[1] + i2 * a.dnb[2] + i3 * a.dnb[3];
    const int64_t i0 = 2 * p;                                  // index of the pair in "cache" order
    if (i0 >= a.n_dims) {                                      // channels past n_dims are copied
        reinterpret_cast<float *>(yd)[i0] = reinterpret_cast<const float *>(xs)[i0];
        reinterpret_cast<float *>(yd)[i0 + 1] = reinterpret_cast<const float *>(xs)[i0 + 1];
        return;
    }
    const float ff = a.freq_factors ? a.freq_factors[p] : 1.0f;
    const float theta_extrap = (float) a.pos[i2] * powf(a.theta_scale, (float) p) / ff;
    // rope_yarn (src/ggml-cpu/ops.cpp:6056-6071)
    float theta = a.freq_scale * theta_extrap, mscale = a.attn_factor;
    if (a.ext_factor != 0.0f) {
        const float yv = ((float) p - a.corr0) / fmaxf(0.001f, a.corr1 - a.corr0);
        const float ramp_mix = (1.0f - fminf(1.0f, fmaxf(0.0f, yv))) * a.ext_factor;
        theta = theta * (1.0f - ramp_mix) + theta_extrap * ramp_mix;
        mscale *= 1.0f + 0.1f * logf(1.0f / a.freq_scale);
    }
    const float c = cosf(theta) * mscale, sn = sinf(theta) * mscale;
    const int64_t e0 = a.neox ? p : i0, e1 = a.neox ? p + a.n_dims / 2 : i0 + 1;
    const float x0 = reinterpret_cast<const float *>(xs)[e0], x1 = reinterpret_cast<const float *>(xs)[e1];
    reinterpret_cast<float *>(yd)[e0] = x0 * c - x1 * sn;
    reinterpret_cast<float *>(yd)[e1] = x0 * sn + x1 * c;
}

// ---------------------------------------------------------------- SOFT_MAX over rows: softmax(x * scale + slope * mask) (+ sinks)
struct softmax_args {
    t4 s;                      // src (dst contiguous, same shape)
    const uint8_t * mask; int mask_f16; int64_t mnb[4]; int64_t mne2, mne3;
    const float * sinks;
    float scale, max_bias, m0, m1; uint32_t n_head_log2;
};
__global__ void __launch_bounds__(256) k_soft_max(const uint8_t * __restrict__ x, float * __restrict__ y, const softmax_args a) {
    __shared__ float red[32];
    const int64_t row = blockIdx.x;
    const int64_t i1 = row % a.s.ne[1], i2 = (row / a.s.ne[1]) % a.s.ne[2], i3 = row / (a.s.ne[1] * a.s.ne[2]);
    const float * xr = reinterpret_cast<const float *>(x + i1 * a.s.nb[1] + i2 * a.s.nb[2] + i3 * a.s.nb[3]);
    float * yr = y + row * a.s.ne[0];
    const uint8_t * mr = a.mask ? a.mask + i1 * a.mnb[1] + (i2 % a.mne2) * a.mnb[2] + (i3 % a.mne3) * a.mnb[3] : nullptr;
    float slope = 1.0f;
    if (a.max_bias > 0.0f) {
        const uint32_t h = (uint32_t) i2;
        slope = h < a.n_head_log2 ? powf(a.m0, (float) (h + 1)) : powf(a.m1, (float) (2 * (h - a.n_head_log2) + 1));
    }
    auto val = [&](int64_t i) -> float {
        float v = xr[i] * a.scale;
        if (mr) v += slope * (a.mask_f16 ? __half2float(reinterpret_cast<const __half *>(mr)[i]) : reinterpret_cast<const float *>(mr)[i]);
        return v;
    };
    float mx = -INFINITY;
    for (int64_t i = threadIdx.x; i < a.s.ne[0]; i += blockDim.x) mx = fmaxf(mx, val(i));
    mx = block_max(mx, red);
    if (a.sinks) mx = fmaxf(mx, a.sinks[i2]);
    float sum = 0.0f;
    for (int64_t i = threadIdx.x; i < a.s.ne[0]; i += blockDim.x) { const float e = expf(val(i) - mx); yr[i] = e; sum += e; }
    sum = block_sum(sum, red);
    if (a.sinks) sum += expf(a.sinks[i2] - mx);
    const float inv = 1.0f / sum;
    for (int64_t i = threadIdx.x; i < a.s.ne[0]; i += blockDim.x) yr[i] *= inv;
}

static bool fill(t4 & t, const int64_t ne[4], const size_t nb[4]) {
    for (int i = 0; i < 4; i++) { if (ne[i] < 0) return false; t.ne[i] = ne[i]; t.nb[i] = (int64_t) nb[i]; }
    return true;
}
static int64_t count(const int64_t ne[4]) { return ne[0] * ne[1] * ne[2] * ne[3]; }

} // namespace b200

using namespace b200;

extern "C" int b200_rms_norm(const float * x_dev, const int64_t ne[4], const size_t nb[4], float * dst_dev, float eps, void * stream) {
    t4 s;
    if (!fill(s, ne, nb) || eps < 0.0f) return B200_ERR_SHAPE;
    if (nb[0] != sizeof(float)) return B200_ERR_STRIDE;
    const int64_t rows = ne[1] * ne[2] * ne[3];
    if (rows == 0 || ne[0] == 0) return B200_OK;
    if (rows > 0x7FFFFFFF) return B200_ERR_SHAPE;
    k_rms_norm<<<(unsigned) rows, ne[0] >= 1024 ? 256 : 64, 0, (cudaStream_t) stream>>>((const uint8_t *) x_dev, s, dst_dev, eps);
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

extern "C" int b200_bin_bcast(int op, const float * a_dev, const int64_t ne_a[4], const size_t nb_a[4],
                              const float * b_dev, const int64_t ne_b[4], const size_t nb_b[4], float * dst_dev, void * stream) {
    t4 sa, sb;
    if ((op != 0 && op != 1) || !fill(sa, ne_a, nb_a) || !fill(sb, ne_b, nb_b)) return B200_ERR_SHAPE;
    for (int i = 0; i < 4; i++) if (ne_b[i] == 0 || ne_a[i] % ne_b[i] != 0) return count(ne_a) == 0 ? B200_OK : B200_ERR_SHAPE;
    const int64_t n = count(ne_a);
    if (n == 0) return B200_OK;
    const unsigned grid = (unsigned) ((n + 255) / 256);
    if (op == 0) k_bin_bcast<0><<<grid, 256, 0, (cudaStream_t) stream>>>((const uint8_t *) a_dev, sa, (const uint8_t *) b_dev, sb, dst_dev, n);
    else         k_bin_bcast<1><<<grid, 256, 0, (cudaStream_t) stream>>>((const uint8_t *) a_dev, sa, (const uint8_t *) b_dev, sb, dst_dev, n);
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

extern "C" int b200_swiglu(const float * x_dev, size_t x_row_stride_bytes, const float * g_dev, size_t g_row_stride_bytes,
                           int64_t nc, int64_t rows, float * dst_dev, void * stream) {
    if (nc < 0 || rows < 0) return B200_ERR_SHAPE;
    if (x_row_stride_bytes % 4 || g_row_stride_bytes % 4) return B200_ERR_STRIDE;
    const int64_t n = nc * rows;
    if (n == 0) return B200_OK;
    k_swiglu<<<(unsigned) ((n + 255) / 256), 256, 0, (cudaStream_t) stream>>>((const uint8_t *) x_dev, (int64_t) x_row_stride_bytes,
                                                                             (const uint8_t *) g_dev, (int64_t) g_row_stride_bytes, dst_dev, nc, n);
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

extern "C" int b200_rope(const float * x_dev, const int64_t ne[4], const size_t nb[4], float * dst_dev, const size_t nb_dst[4],
                         const int32_t * pos_dev, const float * freq_factors_dev, int n_dims, int neox,
                         float freq_base, float freq_scale, float ext_factor, float attn_factor, float corr_dim0, float corr_dim1, void * stream) {
    rope_args a;
    if (!fill(a.s, ne, nb) || n_dims < 0 || n_dims > ne[0] || n_dims % 2 || ne[0] % 2) return B200_ERR_SHAPE;
    if (nb[0] != sizeof(float) || nb_dst[0] != sizeof(float)) return B200_ERR_STRIDE;
    for (int i = 0; i < 4; i++) a.dnb[i] = (int64_t) nb_dst[i];
    a.n_dims = n_dims; a.neox = neox; a.freq_scale = freq_scale; a.ext_factor = ext_factor; a.attn_factor = attn_factor;
    a.theta_scale = powf(freq_base, -2.0f / (float) n_dims);
    a.corr0 = corr_dim0; a.corr1 = corr_dim1; a.pos = pos_dev; a.freq_factors = freq_factors_dev;
    const int64_t n = count(ne) / 2;
    if (n == 0) return B200_OK;
    k_rope<<<(unsigned) ((n + 255) / 256), 256, 0, (cudaStream_t) stream>>>((const uint8_t *) x_dev, (uint8_t *) dst_dev, a);
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

extern "C" int b200_soft_max(const float * x_dev, const int64_t ne[4], const size_t nb[4], const void * mask_dev, int mask_type,
                             const int64_t ne_mask[4], const size_t nb_mask[4], const float * sinks_dev, float scale, float max_bias,
                             float * dst_dev, void * stream) {
    softmax_args a;
    if (!fill(a.s, ne, nb)) return B200_ERR_SHAPE;
    if (nb[0] != sizeof(float)) return B200_ERR_STRIDE;
    a.mask = (const uint8_t *) mask_dev; a.mask_f16 = mask_type == B200_TYPE_F16; a.sinks = sinks_dev;
    a.mne2 = 1; a.mne3 = 1;
    for (int i = 0; i < 4; i++) a.mnb[i] = 0;
    if (mask_dev) {
        if (mask_type != B200_TYPE_F32 && mask_type != B200_TYPE_F16) return B200_ERR_TYPE;
        if (ne_mask[0] != ne[0] || ne_mask[1] < ne[1] || ne_mask[2] < 1 || ne_mask[3] < 1 || ne[2] % ne_mask[2] || ne[3] % ne_mask[3]) return B200_ERR_SHAPE;
        for (int i = 0; i < 4; i++) a.mnb[i] = (int64_t) nb_mask[i];
        a.mne2 = ne_mask[2]; a.mne3 = ne_mask[3];
    }
    a.scale = scale; a.max_bias = max_bias;
    const uint32_t n_head = (uint32_t) ne[2];
    a.n_head_log2 = n_head ? 1u << (uint32_t) floorf(log2f((float) n_head)) : 1u;
    a.m0 = powf(2.0f, -max_bias / (float) a.n_head_log2);
    a.m1 = powf(2.0f, -(max_bias / 2.0f) / (float) a.n_head_log2);
    const int64_t rows = ne[1] * ne[2] * ne[3];
    if (rows == 0 || ne[0] == 0) return B200_OK;
    if (rows > 0x7FFFFFFF) return B200_ERR_SHAPE;
    k_soft_max<<<(unsigned) rows, ne[0] >= 1024 ? 256 : 64, 0, (cudaStream_t) stream>>>((const uint8_t *) x_dev, dst_dev, a);
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}
