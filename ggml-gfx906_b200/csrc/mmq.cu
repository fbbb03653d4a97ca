// mmq.cu — host side of the batched tensor-core MUL_MAT (kernel: mmq_tc.cuh).
// Replaces ggml_cuda_mul_mat_q / mul_mat_q_case / launch_mul_mat_q (reference src/ggml-cuda/mmq.cu:71-143,
// mmq.cuh:3542-3712): converts the activations to bf16 once (instead of quantize_mmq_q8_1), encodes a
// TMA descriptor over them and launches one CTA per 128 x BN output tile.
#include "common.cuh"
#include "mmq_tc.cuh"
#include "mmq_i8.cuh"

#include <mutex>
#include <cstring>

using namespace b200;

namespace {

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_tiled_fn get_encode_tiled() {
    static encode_tiled_fn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) {
            fn = (encode_tiled_fn) p;
        } else {
            cudaGetLastError();
        }
    });
    return fn;
}

template <int T, int BN, bool ALIGNED>
int launch_mmq(const mmq_args & a, const CUtensorMap & tmap, const CUtensorMap & tmap_w, cudaStream_t st) {
    auto kern = k_mmq_tc<T, BN, ALIGNED>;
    const size_t smem = mmq_smem<BN>::total(a.raw_pitch, mmq_rs<T, BN>::value);
    static smem_limit lim;
    if (const int rc = lim.ensure(kern, smem)) return rc;
    const dim3 grid((unsigned) ceil_div(a.m, MMQ_BM), (unsigned) ceil_div(a.n, BN), (unsigned) (a.ksplit > 1 ? a.ksplit : 1));
    kern<<<grid, MMQ_THREADS, smem, st>>>(a, tmap, tmap_w);
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    if (a.ksplit > 1) {
        const dim3 rgrid((unsigned) ceil_div(a.m, 1024), (unsigned) a.n);
        k_mmq_reduce<<<rgrid, 256, 0, st>>>(a.part, a.ksplit, a.m, a.n, a.dst, a.dst_col_stride);
        b200_count_launch(1);
        B200_RETURN_IF_LAUNCH_FAILED();
    }
    return B200_OK;
}

template <int T, int BN>
int launch_mmq_al(mmq_args a, const CUtensorMap & tmap, const CUtensorMap & tmap_w, bool aligned, cudaStream_t st) {
    if (aligned) { a.raw_pitch = slab<T>::pitch; return launch_mmq<T, BN, true>(a, tmap, tmap_w, st); }
    a.raw_pitch = slab<T>::pitch_unaligned;
    return launch_mmq<T, BN, false>(a, tmap, tmap_w, st);
}

template <int T>
int launch_mmq_bn(const mmq_args & a, const CUtensorMap & tmap, const CUtensorMap & tmap_w, int bn, bool aligned, cudaStream_t st) {
    switch (bn) {
        case 32:  return launch_mmq_al<T, 32>(a, tmap, tmap_w, aligned, st);
        case 64:  return launch_mmq_al<T, 64>(a, tmap, tmap_w, aligned, st);
        case 128: return launch_mmq_al<T, 128>(a, tmap, tmap_w, aligned, st);
        case 512: if constexpr (T != T_Q8_0) return launch_mmq_al<T, 512>(a, tmap, tmap_w, aligned, st);     // (Q8_0: pick_bn never says 512)
        default:  return launch_mmq_al<T, 256>(a, tmap, tmap_w, aligned, st);
    }
}

// Split K when the tiles leave more than half of the SMs idle (4096 x k at n = 512: 32-64 tiles on 148 SMs): z CTAs per tile
// contract a share of the 256-deep slabs each and a second small kernel adds their partial tiles in split order.
// Returns the number of splits (1: none) and the slabs per split.
int pick_ksplit(int64_t m, int64_t n, int64_t k, int bn, int sm_count, int * slabs_per_split) {
    const int64_t tiles = ceil_div(m, MMQ_BM) * ceil_div(n, bn), nslab = ceil_div(k, 256);
    *slabs_per_split = (int) nslab;
    static const int max_split = env_int("GGML_B200_MMQ_KSPLIT", 4);
    if (max_split < 2 || bn > 256 || n <= 8 || tiles * 2 > sm_count || nslab < 8) return 1;
    int ks = (int) (sm_count / tiles);
    if (ks > max_split) ks = max_split;
    if (ks > nslab / 4) ks = (int) (nslab / 4);
    if (ks < 2) return 1;
    const int per = (int) ceil_div(nslab, ks);
    ks = (int) ceil_div(nslab, per);            // no empty split
    *slabs_per_split = per;
    return ks < 2 ? 1 : ks;
}

int pick_bn(int type, int64_t m, int64_t n, int64_t k, int sm_count) {
    const int forced = b200_option(B200_OPT_MMQ_BN);
    // 512 columns = two 256-column accumulators per CTA; Q8_0's 272-byte raw slabs do not fit beside two 64 KB activation stages
    const bool can512 = type != T_Q8_0;
    if (forced == 512 && can512) return 512;
    if (forced == 32 || forced == 64 || forced == 128 || forced == 256) return forced;
    if (n <= 32) return 32;
    if (n <= 64) return 64;
    if (n <= 128) return 128;
    // larger n: time = waves x slabs per CTA x the stage time of the tile width, + the partial-tile traffic of a K split.
    // Stage times in units of the MMA time of one column (measured, profiles/r01g_prefill_probe.txt, r03j_prefill_bn512.txt):
    // the weight expansion costs ~600 whatever the width, the MMAs their columns; a 512-column tile (two accumulators fed by
    // one expansion) measured 1.76 x the 256-column stage (shared-memory bandwidth: A is read once per accumulator).
    const int widths[3] = { 128, 256, 512 };
    const double stage[3] = { 600 + 128, 600 + 256, 1507 };
    int best = 256; double best_cost = 0;
    for (int i = 0; i < 3; i++) {
        const int bn = widths[i];
        if (bn == 512 && !(can512 && n > 256)) continue;
        int per = 0;
        const int ks = pick_ksplit(m, n, k, bn, sm_count, &per);
        const int64_t ctas = ceil_div(m, MMQ_BM) * ceil_div(n, bn) * ks;
        double cost = (double) ceil_div(ctas, sm_count) * per * 4.0 * stage[i];
        if (ks > 1) cost += (double) (ks + 1) * (double) m * (double) n * 1.0e-3;      // partial tiles written and re-read (~6 TB/s)
        if (best_cost == 0 || cost < best_cost * 0.999) { best = bn; best_cost = cost; }
    }
    return best;
}

} // namespace

// ---------------------------------------------------------------- the int8 route (mmq_i8.cuh)
static size_t i8_yq_bytes(int64_t k, int64_t n_pad) { return ((size_t) n_pad * (size_t) k + 255) & ~(size_t) 255; }
static size_t i8_ds_bytes(int64_t k, int64_t n_pad) { return (size_t) ((k + 127) / 128) * (size_t) n_pad * 4 * sizeof(float2); }

template <int T, int BN>
static int launch_mmq_i8(const mmq_i8_args & a, const CUtensorMap & tmap_b, const CUtensorMap & tmap_w, cudaStream_t st) {
    auto kern = k_mmq_i8<T, BN>;
    const size_t smem = mmq_i8_smem<BN>::total(a.raw_pitch, MMQI_RS);
    static smem_limit lim;
    if (const int rc = lim.ensure(kern, smem)) return rc;
    const dim3 grid((unsigned) ceil_div(a.m, MMQI_BM), (unsigned) ceil_div(a.n, BN));
    kern<<<grid, MMQI_THREADS, smem, st>>>(a, tmap_b, tmap_w);
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}
template <int T>
static int launch_mmq_i8_bn(mmq_i8_args a, const CUtensorMap & tmap_b, const CUtensorMap & tmap_w, int bn, cudaStream_t st) {
    a.raw_pitch = slab<T>::pitch;
    switch (bn) {
        case 32:  return launch_mmq_i8<T, 32>(a, tmap_b, tmap_w, st);
        case 64:  return launch_mmq_i8<T, 64>(a, tmap_b, tmap_w, st);
        default:  return launch_mmq_i8<T, 128>(a, tmap_b, tmap_w, st);
    }
}

// B200_NOT_APPLICABLE: this format / alignment has no int8 route (the caller takes the bf16 kernel)
static int mul_mat_q_i8(int type, const void * w_dev, size_t w_row_stride_bytes, int64_t m, int64_t k,
                        const float * x_dev, size_t x_row_stride_bytes, int64_t n, float * dst_dev, size_t dst_col_stride_bytes,
                        void * workspace_dev, cudaStream_t st) {
    if (type == T_Q6_K) return B200_NOT_APPLICABLE;                    // scales per 16 weights: two groups per K = 32 MMA
    if ((uintptr_t) w_dev % 16 != 0 || w_row_stride_bytes % 16 != 0 || b200_row_bytes(type, k) % 4 != 0 || k % 16 != 0) return B200_NOT_APPLICABLE;
    encode_tiled_fn encode = get_encode_tiled();
    if (!encode) return B200_ERR_NO_DEVICE;
    const int forced = b200_option(B200_OPT_MMQ_BN);
    const int bn = (forced == 32 || forced == 64 || forced == 128) ? forced : (n <= 32 ? 32 : (n <= 64 ? 64 : 128));
    const int64_t n_pad = ceil_div(n, 128) * 128;
    int8_t * yq = (int8_t *) (((uintptr_t) workspace_dev + 255) & ~(uintptr_t) 255);
    float2 * ds = (float2 *) ((uint8_t *) yq + i8_yq_bytes(k, n_pad));
    // 1. activations -> q8_1 (int8 plane + (d, s) pairs), the analogue of quantize_mmq_q8_1 (src/ggml-cuda/quantize.cu:50-143)
    for (int64_t r0 = 0; r0 < n_pad; r0 += 65535) {
        const int64_t nr = n_pad - r0 < 65535 ? n_pad - r0 : 65535;
        const dim3 grid((unsigned) ceil_div(ceil_div(k, 128) * 4, 8), (unsigned) nr);
        k_quantize_mmq_q8_1<<<grid, 256, 0, st>>>((const uint8_t *) x_dev + r0 * x_row_stride_bytes, (int64_t) x_row_stride_bytes, k,
                                                  n - r0 > 0 ? n - r0 : 0, n_pad, yq + r0 * k, ds + r0 * 4);
        b200_count_launch(1);
    }
    B200_RETURN_IF_LAUNCH_FAILED();
    // 2. descriptors: activations {k, n_pad} bytes, box {128, bn}, 128-byte swizzle; raw weights as in the bf16 route
    CUtensorMap tmap_b, tmap_w;
    const cuuint32_t estr[2] = { 1, 1 };
    {
        const cuuint64_t gdim[2] = { (cuuint64_t) k, (cuuint64_t) n_pad };
        const cuuint64_t gstride[1] = { (cuuint64_t) k };
        const cuuint32_t box[2] = { 128, (cuuint32_t) bn };
        if (encode(&tmap_b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, yq, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return B200_NOT_APPLICABLE;
        const int slab_words = (type == T_Q8_0 ? 272 : (type == T_Q5_K ? 176 : 144)) / 4;
        const cuuint64_t wdim[2] = { (cuuint64_t) (b200_row_bytes(type, k) / 4), (cuuint64_t) m };
        const cuuint64_t wstride[1] = { (cuuint64_t) w_row_stride_bytes };
        const cuuint32_t wbox[2] = { (cuuint32_t) slab_words, (cuuint32_t) MMQI_BM };
        if (encode(&tmap_w, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void *>(w_dev), wdim, wstride, wbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return B200_NOT_APPLICABLE;
    }
    mmq_i8_args a;
    a.m = m; a.k = k; a.n = n; a.n_pad = n_pad; a.dst = dst_dev; a.dst_col_stride = (int64_t) (dst_col_stride_bytes / 4);
    a.ds = ds; a.nblk = (int) (k / type_qk(type)); a.raw_pitch = 0;
    switch (type) {
        case T_Q4_0: return launch_mmq_i8_bn<T_Q4_0>(a, tmap_b, tmap_w, bn, st);
        case T_Q8_0: return launch_mmq_i8_bn<T_Q8_0>(a, tmap_b, tmap_w, bn, st);
        case T_Q4_K: return launch_mmq_i8_bn<T_Q4_K>(a, tmap_b, tmap_w, bn, st);
        case T_Q5_K: return launch_mmq_i8_bn<T_Q5_K>(a, tmap_b, tmap_w, bn, st);
        default: return B200_NOT_APPLICABLE;
    }
}

extern "C" size_t b200_mul_mat_q_workspace_size(int64_t m, int64_t k, int64_t n) {
    (void) m;
    if (k <= 0 || n <= 0) return 256;
    const size_t bf16 = (size_t) n * (size_t) k * sizeof(__nv_bfloat16);
    const int64_t n_pad = ceil_div(n, 128) * 128;
    const size_t i8 = i8_yq_bytes(k, n_pad) + i8_ds_bytes(k, n_pad);
    // split-K partial tiles (at most 4 splits, and only for matrices of at most sm_count / 2 tiles: a few MB)
    int sm = 0, dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sm <= 0) { cudaGetLastError(); sm = 148; }
    size_t part = 0;
    for (int bn : { 32, 64, 128, 256, 512 }) {
        int per = 0;
        const int ks = pick_ksplit(m, n, k, bn, sm, &per);
        if (ks > 1) { const size_t p = (size_t) ks * (size_t) m * (size_t) n * sizeof(float) + 256; if (p > part) part = p; }
    }
    return (((bf16 > i8 ? bf16 : i8) + 255) & ~(size_t) 255) + part + 512;
}

extern "C" int b200_mul_mat_q(int type, const void * w_dev, size_t w_row_stride_bytes, int64_t m, int64_t k,
                              const float * x_dev, size_t x_row_stride_bytes, int64_t n,
                              float * dst_dev, size_t dst_col_stride_bytes,
                              void * workspace_dev, size_t workspace_bytes, void * stream) {
    if (!type_is_quant(type)) return B200_ERR_TYPE;
    if (m < 0 || k < 0 || n < 0 || k % type_qk(type) != 0) return B200_ERR_SHAPE;
    if (dst_col_stride_bytes % 4 != 0 || w_row_stride_bytes % 2 != 0 || x_row_stride_bytes % 4 != 0) return B200_ERR_STRIDE;
    if (m == 0 || n == 0) return B200_OK;
    cudaStream_t st = (cudaStream_t) stream;
    if (k == 0) return -(int) cudaMemset2DAsync(dst_dev, dst_col_stride_bytes, 0, (size_t) m * 4, (size_t) n, st);
    if (!workspace_dev || workspace_bytes < b200_mul_mat_q_workspace_size(m, k, n)) return B200_ERR_WORKSPACE;
    if (k > 0x7FFFFFFF || n > 0x7FFFFFFF) return B200_ERR_SHAPE;

    // The q8_1 int8 route (option mmq_route = 2; measured slower than the bf16 route on B200, DESIGN.md §4.4: the per-block
    // drain of the accumulators is bound by the CUDA cores, not by the tensor pipe) ..
    if (b200_option(B200_OPT_MMQ_ROUTE) == 2) {
        const int rc = mul_mat_q_i8(type, w_dev, w_row_stride_bytes, m, k, x_dev, x_row_stride_bytes, n, dst_dev, dst_col_stride_bytes, workspace_dev, st);
        if (rc != B200_NOT_APPLICABLE) return rc;
    }
    // .. and the bf16 route.
    // 1. activations -> bf16 [n][k]
    uint16_t * xb = (uint16_t *) (((uintptr_t) workspace_dev + 255) & ~(uintptr_t) 255);
    for (int64_t r0 = 0; r0 < n; r0 += 65535) {
        const int64_t nr = n - r0 < 65535 ? n - r0 : 65535;
        const dim3 grid((unsigned) ceil_div(k, 1024), (unsigned) nr);
        k_f32_to_bf16<<<grid, 256, 0, st>>>((const uint8_t *) x_dev + r0 * x_row_stride_bytes, (int64_t) x_row_stride_bytes, k, xb + r0 * k);
        b200_count_launch(1);
    }
    B200_RETURN_IF_LAUNCH_FAILED();

    // 2. TMA descriptor over the bf16 activations: dims {k, n}, box {64, BN}, 128-byte swizzle, zero fill
    const dev_caps & caps = current_caps();
    const int bn = pick_bn(type, m, n, k, caps.sm_count);
    encode_tiled_fn encode = get_encode_tiled();
    if (!encode) return B200_ERR_NO_DEVICE;
    CUtensorMap tmap;
    const cuuint64_t gdim[2]    = { (cuuint64_t) k, (cuuint64_t) n };
    const cuuint64_t gstride[1] = { (cuuint64_t) k * sizeof(__nv_bfloat16) };
    const cuuint32_t box[2]     = { (cuuint32_t) MMQ_BK, (cuuint32_t) (bn > 256 ? 256 : bn) };   // (512: two boxes per stage)
    const cuuint32_t estr[2]    = { 1, 1 };
    const CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, xb, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) return B200_ERR_STRIDE;

    // 3. the tile kernel
    mmq_args a;
    a.w = (const uint8_t *) w_dev; a.w_row_stride = (int64_t) w_row_stride_bytes; a.m = m; a.k = k; a.n = n;
    a.dst = dst_dev; a.dst_col_stride = (int64_t) (dst_col_stride_bytes / 4);
    a.nblk = (int) (k / type_qk(type)); a.raw_pitch = 0;
    a.ksplit = pick_ksplit(m, n, k, bn, caps.sm_count, &a.slabs_per_split);
    a.part = (float *) (((uintptr_t) (xb + (size_t) n * (size_t) k) + 255) & ~(uintptr_t) 255);      // behind the bf16 activations
    const bool aligned0 = ((uintptr_t) w_dev % 16 == 0) && (w_row_stride_bytes % 16 == 0) &&
                         !b200_option(B200_OPT_MMQ_FORCE_UNALIGNED);
    // aligned rows: a second descriptor over the raw weight bytes as 32-bit words, dims {row words, m}, box {slab words, 128}
    bool aligned = aligned0;
    CUtensorMap tmap_w;
    memset(&tmap_w, 0, sizeof(tmap_w));
    if (aligned && b200_row_bytes(type, k) % 4 != 0) aligned = false;      // the descriptor counts the row in 32-bit words
    if (aligned) {
        // (Q6_K: 60-word box over 210-byte blocks, the expander removes the 2-byte skew of odd slabs: slab_tma<> in mmq_tc.cuh)
        const int slab_words = type == T_Q6_K ? 60 : (type == T_Q8_0 ? 272 : (type == T_Q5_K ? 176 : 144)) / 4;
        const cuuint64_t wdim[2]    = { (cuuint64_t) (b200_row_bytes(type, k) / 4), (cuuint64_t) m };
        const cuuint64_t wstride[1] = { (cuuint64_t) w_row_stride_bytes };
        const cuuint32_t wbox[2]    = { (cuuint32_t) slab_words, (cuuint32_t) MMQ_BM };
        const CUresult wr = encode(&tmap_w, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void *>(w_dev), wdim, wstride, wbox, estr,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (wr != CUDA_SUCCESS) aligned = false;          // odd geometry: per-row copies
    }
    switch (type) {
        case T_Q4_0: return launch_mmq_bn<T_Q4_0>(a, tmap, tmap_w, bn, aligned, st);
        case T_Q8_0: return launch_mmq_bn<T_Q8_0>(a, tmap, tmap_w, bn, aligned, st);
        case T_Q4_K: return launch_mmq_bn<T_Q4_K>(a, tmap, tmap_w, bn, aligned, st);
        case T_Q5_K: return launch_mmq_bn<T_Q5_K>(a, tmap, tmap_w, bn, aligned, st);
        case T_Q6_K: return launch_mmq_bn<T_Q6_K>(a, tmap, tmap_w, bn, aligned, st);
        default: return B200_ERR_TYPE;
    }
}
