// quantize.cu — f32 activations -> q8_1 (32 values: d = amax/127, q = round(x/d), s = d*sum(q)).
//
// Value contract: quantize_row_q8_1_ref (reference src/ggml-quants.c:225-258), bit-exact:
//   d  = amax / 127.0f            (IEEE division)
//   id = d ? 1.0f / d : 0.0f      (IEEE division)
//   q  = roundf(x * id)           (half away from zero)
//   block.d = fp16(d), block.s = fp16((float)sum(q) * d)     (d unrounded f32)
// Two output forms share the arithmetic:
//   * block_q8_1 (36 B) — the public, checkable form (b200_quantize_row_q8_1);
//   * planar scratch for the matvec: int8 plane [n][k] (16-byte loadable) + f32 plane of the
//     fp16-rounded d per 32 (b200_quantize_acts).  s is not stored: the matvec takes integer sums.
//
// One warp per 32-value block (lane = element): max by shuffle, sum by shuffle.  Activation
// quantization is latency-, not bandwidth-bound (k=4096: 16 KiB in, 4.6 KiB out).
#include "common.cuh"

namespace b200 {

struct q8_result { int q; float d; int sum; };

__device__ __forceinline__ q8_result quantize32(float x) {
    const float amax = warp_max(fabsf(x));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    // float -> int8 as the reference's x86 build does it (cvttss2si, then the low byte): NaN and
    // |v| >= 2^31 give INT_MIN, whose low byte is 0; in-range values wrap.  Only reachable when
    // 1/d overflows (f32-denormal inputs) — the golden vectors from the reference pin it.
    const float v  = roundf(__fmul_rn(x, id));
    const int   qi = fabsf(v) < 2147483648.0f ? (int) v : (int) 0x80000000;
    const int   q  = (int) (int8_t) (qi & 0xFF);
    int sum = q;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
    return { q, d, sum };
}

// 8 lanes per 32-value block (a float4 each), FOUR blocks per lane group: grid (ceil(k/32 / 128), n), block 256 = 128 blocks
// of 32 values; a thread's four 16-byte loads are issued before any arithmetic (64 bytes in flight per thread: with one
// block per group the kernel sat at 48 % of the HBM peak, latency-bound).  Same arithmetic as quantize32 (max and integer sum
// are order-independent); the one-value-per-lane form measured 1.25 TB/s (profiles/r01i_support_probe.txt).
__global__ void __launch_bounds__(256) k_quantize_q8_1(const uint8_t * __restrict__ x, int64_t x_row_stride, int64_t k,
                                                       uint8_t * __restrict__ y) {
    const int l8 = threadIdx.x & 7;
    const int64_t row = blockIdx.y;
    const float * xr = reinterpret_cast<const float *>(x + row * x_row_stride);
    const bool al = (reinterpret_cast<uintptr_t>(xr) & 15) == 0;
    float v[4][4];
    int64_t kbs[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int64_t kb = (int64_t) blockIdx.x * 128 + u * 32 + (threadIdx.x >> 3);
        kbs[u] = kb;
        v[u][0] = v[u][1] = v[u][2] = v[u][3] = 0.0f;
        if (kb * 32 < k) {                          // whole 8-lane groups are live or not: shuffles stay inside a group
            if (al) {
                const float4 t = __ldg(reinterpret_cast<const float4 *>(xr + kb * 32 + l8 * 4));
                v[u][0] = t.x; v[u][1] = t.y; v[u][2] = t.z; v[u][3] = t.w;
            } else {
#pragma unroll
                for (int j = 0; j < 4; j++) v[u][j] = xr[kb * 32 + l8 * 4 + j];
            }
        }
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
        float amax = fmaxf(fmaxf(fabsf(v[u][0]), fabsf(v[u][1])), fmaxf(fabsf(v[u][2]), fabsf(v[u][3])));
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, o));
        const float d  = __fdiv_rn(amax, 127.0f);
        const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
        uint32_t packed = 0;
        int sum = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float r  = roundf(__fmul_rn(v[u][j], id));
            const int   qi = fabsf(r) < 2147483648.0f ? (int) r : (int) 0x80000000;     // see quantize32
            const int   q  = (int) (int8_t) (qi & 0xFF);
            packed |= (uint32_t) (q & 0xFF) << (8 * j);
            sum += q;
        }
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
        if (kbs[u] * 32 >= k) continue;
        uint8_t * b = y + (row * (k / 32) + kbs[u]) * 36;                                // blocks are 4-byte aligned (36 B)
        *reinterpret_cast<uint32_t *>(b + 4 + 4 * l8) = packed;
        if (l8 == 0) {
            const __half hd = __float2half_rn(d);
            const __half hs = __float2half_rn(__fmul_rn((float) sum, d));
            *reinterpret_cast<__half2 *>(b) = __halves2half2(hd, hs);
        }
    }
}

__global__ void __launch_bounds__(256) k_quantize_acts(const uint8_t * __restrict__ x, int64_t x_row_stride, int64_t k, int64_t n,
                                                       int8_t * __restrict__ q, float * __restrict__ dd) {
    const int64_t kb = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (kb * 32 >= k) return;
    const int64_t row = blockIdx.y;
    const float v = reinterpret_cast<const float *>(x + row * x_row_stride)[kb * 32 + lane];
    const q8_result r = quantize32(v);
    q[row * k + kb * 32 + lane] = (int8_t) r.q;
    if (lane == 0) dd[row * (k / 32) + kb] = __half2float(__float2half_rn(r.d));
}

// ---------------------------------------------------------------- f32 -> Q8_0 / Q4_0 (CPY f32 -> quant)
// replaces cpy_f32_q<cpy_blck_f32_q8_0 / _q4_0> (reference src/ggml-cuda/cpy.cu:66-92, cpy-utils.cuh) with the value contract of
// the CPU definitions quantize_row_q8_0_ref / quantize_row_q4_0_ref (src/ggml-quants.c:199-222, 36-71), bit-exact:
//   Q8_0: d = amax / 127, q = roundf(x / d)                         Q4_0: d = (the FIRST value of largest magnitude) / -8,
//                                                                         q = min(15, (int8) (x / d + 8.5f))
// One warp per 32-value block (lane = element); rows of k values x_row_stride bytes apart, blocks of a row contiguous.
// rows (i1, i2, i3) of a 4-D tensor: strides in bytes, dim 0 dense on both sides (ggml CPY between tensors of one shape)
struct qrows_geo { int64_t k, ne1, ne2; int64_t xs[3], ys[3]; };
template <int T>
__global__ void __launch_bounds__(256) k_quantize_blocks(const uint8_t * __restrict__ x, uint8_t * __restrict__ y, const qrows_geo g, int64_t row0) {
    const int64_t kb = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (kb * 32 >= g.k) return;
    const int64_t row = row0 + blockIdx.y, i1 = row % g.ne1, i2 = (row / g.ne1) % g.ne2, i3 = row / (g.ne1 * g.ne2);
    const float v = reinterpret_cast<const float *>(x + i1 * g.xs[0] + i2 * g.xs[1] + i3 * g.xs[2])[kb * 32 + lane];
    uint8_t * b = y + i1 * g.ys[0] + i2 * g.ys[1] + i3 * g.ys[2] + kb * fmt<T>::bytes;
    if (T == T_Q8_0) {
        const float amax = warp_max(fabsf(v));
        const float d  = __fdiv_rn(amax, 127.0f);
        const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
        b[2 + lane] = (uint8_t) (int8_t) roundf(__fmul_rn(v, id));
        if (lane == 0) *reinterpret_cast<__half *>(b) = __float2half_rn(d);
    } else {
        // the first element of largest |v| (the reference's loop keeps the earlier one on ties): key = |v| bits : 31 - lane
        unsigned long long key = ((unsigned long long) (__float_as_uint(v) & 0x7FFFFFFFu) << 5) | (unsigned) (31 - lane);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { const unsigned long long t = __shfl_xor_sync(0xFFFFFFFFu, key, o); key = t > key ? t : key; }
        const float mx = (key >> 5) != 0ull ? __shfl_sync(0xFFFFFFFFu, v, 31 - (int) (key & 31u)) : 0.0f;
        const float d  = __fdiv_rn(mx, -8.0f);
        const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
        const int q = min(15, (int) (int8_t) (int) __fadd_rn(__fmul_rn(v, id), 8.5f));       // (int8_t) of a float in [0.5, 16.5]: truncation
        const int hi = __shfl_down_sync(0xFFFFFFFFu, q, 16);
        if (lane < 16) b[2 + lane] = (uint8_t) ((q & 0xFF) | (hi << 4));
        if (lane == 0) *reinterpret_cast<__half *>(b) = __float2half_rn(d);
    }
}

} // namespace b200

using namespace b200;

int b200::b200_quantize_acts_planes(const float * x_dev, size_t x_row_stride_bytes, int64_t k, int64_t n,
                                    int8_t * q, float * dd, cudaStream_t st) {
    for (int64_t r0 = 0; r0 < n; r0 += 65535) {
        const int64_t nr = n - r0 < 65535 ? n - r0 : 65535;
        const dim3 grid((unsigned) ceil_div(k / 32, 8), (unsigned) nr);
        k_quantize_acts<<<grid, 256, 0, st>>>((const uint8_t *) x_dev + r0 * x_row_stride_bytes, (int64_t) x_row_stride_bytes, k, n,
                                              q + r0 * k, dd + r0 * (k / 32));
        b200_count_launch(1);
    }
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

extern "C" int b200_quantize_row_q8_1(const float * x_dev, size_t x_row_stride_bytes, int64_t k, int64_t n, void * y_dev, void * stream) {
    if (k < 0 || n < 0 || k % 32 != 0) return B200_ERR_SHAPE;
    if (x_row_stride_bytes % 4 != 0) return B200_ERR_STRIDE;
    if (k == 0 || n == 0) return B200_OK;
    for (int64_t r0 = 0; r0 < n; r0 += 65535) {
        const int64_t nr = n - r0 < 65535 ? n - r0 : 65535;
        const dim3 grid((unsigned) ceil_div(k / 32, 128), (unsigned) nr);
        k_quantize_q8_1<<<grid, 256, 0, (cudaStream_t) stream>>>((const uint8_t *) x_dev + r0 * x_row_stride_bytes, (int64_t) x_row_stride_bytes, k,
                                                                (uint8_t *) y_dev + r0 * (k / 32) * 36);
        b200_count_launch(1);
    }
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

extern "C" size_t b200_acts_scratch_bytes(int64_t k, int64_t n) {
    if (k <= 0 || n <= 0) return 0;
    const size_t qbytes = ((size_t) n * (size_t) k + 255) / 256 * 256;
    return qbytes + (size_t) n * (size_t) (k / 32) * sizeof(float);
}

extern "C" int b200_quantize_acts(const float * x_dev, size_t x_row_stride_bytes, int64_t k, int64_t n, void * scratch_dev, void * stream) {
    if (k < 0 || n < 0 || k % 32 != 0) return B200_ERR_SHAPE;
    if (x_row_stride_bytes % 4 != 0) return B200_ERR_STRIDE;
    if (k == 0 || n == 0) return B200_OK;
    if (!scratch_dev) return B200_ERR_WORKSPACE;
    const size_t qbytes = ((size_t) n * (size_t) k + 255) / 256 * 256;
    int8_t * q  = (int8_t *) scratch_dev;
    float  * dd = (float *) ((uint8_t *) scratch_dev + qbytes);
    return b200_quantize_acts_planes(x_dev, x_row_stride_bytes, k, n, q, dd, (cudaStream_t) stream);
}

// rows of f32 -> rows of Q8_0 / Q4_0 blocks (b200_quantize_rows / _4d, include/ggml_b200.h)
static int quantize_rows_launch(int type, const void * x_dev, void * y_dev, const qrows_geo & g, int64_t n, cudaStream_t st) {
    for (int64_t r0 = 0; r0 < n; r0 += 65535) {
        const int64_t nr = n - r0 < 65535 ? n - r0 : 65535;
        const dim3 grid((unsigned) ceil_div(g.k / 32, 8), (unsigned) nr);
        if (type == T_Q8_0) k_quantize_blocks<T_Q8_0><<<grid, 256, 0, st>>>((const uint8_t *) x_dev, (uint8_t *) y_dev, g, r0);
        else                k_quantize_blocks<T_Q4_0><<<grid, 256, 0, st>>>((const uint8_t *) x_dev, (uint8_t *) y_dev, g, r0);
        b200_count_launch(1);
    }
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}
extern "C" int b200_quantize_rows(int type, const float * x_dev, size_t x_row_stride_bytes, int64_t k, int64_t n,
                                  void * y_dev, size_t y_row_stride_bytes, void * stream) {
    if (type != T_Q8_0 && type != T_Q4_0) return B200_ERR_TYPE;
    if (k < 0 || n < 0 || k % 32 != 0) return B200_ERR_SHAPE;
    if (x_row_stride_bytes % 4 != 0 || y_row_stride_bytes % 2 != 0 || ((uintptr_t) y_dev & 1)) return B200_ERR_STRIDE;
    if (k == 0 || n == 0) return B200_OK;
    qrows_geo g;
    g.k = k; g.ne1 = n; g.ne2 = 1;
    g.xs[0] = (int64_t) x_row_stride_bytes; g.xs[1] = g.xs[2] = 0; g.ys[0] = (int64_t) y_row_stride_bytes; g.ys[1] = g.ys[2] = 0;
    return quantize_rows_launch(type, x_dev, y_dev, g, n, (cudaStream_t) stream);
}
extern "C" int b200_quantize_rows_4d(int type, const float * x_dev, const int64_t ne[4], const size_t nb_x[4],
                                     void * y_dev, const size_t nb_y[4], void * stream) {
    if (type != T_Q8_0 && type != T_Q4_0) return B200_ERR_TYPE;
    if (ne[0] < 0 || ne[1] < 0 || ne[2] < 0 || ne[3] < 0 || ne[0] % 32 != 0) return B200_ERR_SHAPE;
    if (nb_x[0] != sizeof(float) || nb_y[0] != (size_t) type_bytes(type) || ((uintptr_t) x_dev & 3) || ((uintptr_t) y_dev & 1)) return B200_ERR_STRIDE;
    for (int d = 1; d < 4; d++) if (nb_x[d] % 4 != 0 || nb_y[d] % 2 != 0) return B200_ERR_STRIDE;
    const int64_t n = ne[1] * ne[2] * ne[3];
    if (ne[0] == 0 || n == 0) return B200_OK;
    qrows_geo g;
    g.k = ne[0]; g.ne1 = ne[1]; g.ne2 = ne[2];
    for (int d = 0; d < 3; d++) { g.xs[d] = (int64_t) nb_x[d + 1]; g.ys[d] = (int64_t) nb_y[d + 1]; }
    return quantize_rows_launch(type, x_dev, y_dev, g, n, (cudaStream_t) stream);
}
