// mmvq.cu — decode-time quantized mat-vec: dst[m, n] = W[m, k] x X[n, k]^T for n <= 8, weights in a
// ggml block format, activations pre-quantized to q8 (quantize.cu).  HBM-bound: the weights are read
// exactly once.  Replaces mul_mat_vec_q + launchers (reference src/ggml-cuda/mmvq.cu:139-343, 500-570).
//
// Two kernels:
//   k_mmvq_generic — one warp per weight row, lanes stride over 32-weight sub-blocks, any alignment,
//                    any strides.  The always-correct path (tiny / ragged / permuted cases).
//   k_mmvq_stream  — (mmvq_stream.cuh) TMA-staged K-stationary streaming kernel for full-size matrices.
#include "common.cuh"
#include "unpack.cuh"
#include "mmvq_stream.cuh"
#include "mmvq_fused.cuh"
#include "mmvq_batch.cuh"
#include "mmvq_cols.cuh"

#include <cstring>
#include <vector>

namespace b200 {

struct mmvq_args {
    const uint8_t * w;      // weights
    const int8_t  * yq;     // activation int8 plane  [cols_total][k]
    const float   * yd;     // activation scale plane [cols_total][k/32]
    float         * dst;
    int64_t m, k;
    int64_t w_row_stride;   // bytes
    int64_t w_nb2, w_nb3;   // bytes, weight channel strides
    int64_t dst_col_stride; // floats between consecutive columns of dst
    int64_t dst_nb2, dst_nb3; // floats
    int64_t ne12, ne13;     // activation/dst channels (grid y, z)
    int64_t r2, r3;         // broadcast ratios ne12/ne02, ne13/ne03
    int64_t n;              // columns per channel (of this launch)
    int64_t col_stride_n;   // columns per channel in the scratch (0: = n)
    // MUL_MAT_ID (mixture of experts): channel (i12 = slot e, i13 = token t) multiplies expert ids[e, t] of the stack
    // a[k, m, n_as] with activation row (e % ne11, t); NULL for plain MUL_MAT
    const uint8_t * ids;
    int64_t ids_nb0, ids_nb1, n_as, ne11;
    unsigned * bad;         // sticky error word: an expert id outside [0, n_as)
};

template <int T, int NCOLS>
__global__ void __launch_bounds__(128) k_mmvq_generic(const mmvq_args a) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row = (int64_t) blockIdx.x * 4 + warp;
    if (row >= a.m) return;
    const int64_t i12 = blockIdx.y, i13 = blockIdx.z;
    const uint8_t * wrow;
    int64_t col0;                                             // first column of this channel in the scratch
    if (a.ids) {
        const int32_t e = *reinterpret_cast<const int32_t *>(a.ids + i12 * a.ids_nb0 + i13 * a.ids_nb1);
        if (e < 0 || e >= a.n_as) {                           // the reference asserts (src/ggml-cpu/ggml-cpu.c:1556): flag it, write zeros
            if (lane == 0) { if (a.bad && row == 0) atomicOr_system(a.bad, 1u); a.dst[i12 * a.dst_nb2 + i13 * a.dst_nb3 + row] = 0.0f; }
            return;
        }
        wrow = a.w + row * a.w_row_stride + (int64_t) e * a.w_nb2;
        col0 = i13 * a.ne11 + i12 % a.ne11;
    } else {
        wrow = a.w + row * a.w_row_stride + (i12 / a.r2) * a.w_nb2 + (i13 / a.r3) * a.w_nb3;
        col0 = (i13 * a.ne12 + i12) * (a.col_stride_n ? a.col_stride_n : a.n);
    }
    const int64_t nsub = a.k / 32;

    float acc[NCOLS];
#pragma unroll
    for (int c = 0; c < NCOLS; c++) acc[c] = 0.0f;

    for (int64_t u = lane; u < nsub; u += 32) {
        constexpr int S = subs<T>::n;
        const int64_t blk = u / S;
        sub32 w;
        unpack32<T, ld_global_a2>::run(wrow + blk * fmt<T>::bytes, (int) (u - blk * S), w);
#pragma unroll
        for (int c = 0; c < NCOLS; c++) {
            const int4 * yp = reinterpret_cast<const int4 *>(a.yq + (col0 + c) * a.k + u * 32);
            const int4 y0 = __ldg(yp), y1 = __ldg(yp + 1);
            const int y[8] = { y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w };
            acc[c] += dot_sub32(w, y, __ldg(a.yd + (col0 + c) * nsub + u));
        }
    }
    float * out = a.dst + i12 * a.dst_nb2 + i13 * a.dst_nb3 + row;
#pragma unroll
    for (int c = 0; c < NCOLS; c++) {
        const float v = warp_sum(acc[c]);
        if (lane == 0) out[c * a.dst_col_stride] = v;
    }
}

template <int T>
static int launch_generic(const mmvq_args & a, cudaStream_t st) {
    const dim3 grid((unsigned) ceil_div(a.m, 4), (unsigned) a.ne12, (unsigned) a.ne13);
    switch (a.n) {
        case 1: k_mmvq_generic<T, 1><<<grid, 128, 0, st>>>(a); break;
        case 2: k_mmvq_generic<T, 2><<<grid, 128, 0, st>>>(a); break;
        case 3: k_mmvq_generic<T, 3><<<grid, 128, 0, st>>>(a); break;
        case 4: k_mmvq_generic<T, 4><<<grid, 128, 0, st>>>(a); break;
        case 5: k_mmvq_generic<T, 5><<<grid, 128, 0, st>>>(a); break;
        case 6: k_mmvq_generic<T, 6><<<grid, 128, 0, st>>>(a); break;
        case 7: k_mmvq_generic<T, 7><<<grid, 128, 0, st>>>(a); break;
        case 8: k_mmvq_generic<T, 8><<<grid, 128, 0, st>>>(a); break;
        default: return B200_ERR_SHAPE;
    }
    return B200_OK;
}

static int mmvq_dispatch(int type, const mmvq_args & a, cudaStream_t st) {
    if (a.m == 0 || a.n == 0 || a.ne12 == 0 || a.ne13 == 0) return B200_OK;
    if (a.ne12 > 65535 || a.ne13 > 65535) return B200_ERR_SHAPE;
    int rc = B200_ERR_TYPE;
    // fast path: plain 2-D matrix, rows contiguous (the five hot-path formats)
    if (a.ne12 == 1 && a.ne13 == 1 && type_is_quant(type)) {
        rc = mmvq_stream_try(type, a.w, a.w_row_stride, a.m, a.k, a.yq, a.yd, a.n, a.dst, a.dst_col_stride, st);
        if (rc == B200_OK) return rc;
        if (rc != B200_STREAM_NOT_APPLICABLE) return rc;
    }
    switch (type) {
        case T_Q4_0: rc = launch_generic<T_Q4_0>(a, st); break;
        case T_Q8_0: rc = launch_generic<T_Q8_0>(a, st); break;
        case T_Q4_K: rc = launch_generic<T_Q4_K>(a, st); break;
        case T_Q5_K: rc = launch_generic<T_Q5_K>(a, st); break;
        case T_Q6_K: rc = launch_generic<T_Q6_K>(a, st); break;
        case T_Q4_1: rc = launch_generic<T_Q4_1>(a, st); break;
        case T_Q5_0: rc = launch_generic<T_Q5_0>(a, st); break;
        case T_Q5_1: rc = launch_generic<T_Q5_1>(a, st); break;
        case T_IQ4_NL: rc = launch_generic<T_IQ4_NL>(a, st); break;
        default: return B200_ERR_TYPE;
    }
    if (rc != B200_OK) return rc;
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

} // namespace b200

using namespace b200;

extern "C" int b200_mul_mat_vec_q(int type, const void * w_dev, size_t w_row_stride_bytes, int64_t m, int64_t k,
                                  const void * acts_scratch_dev, int64_t n,
                                  float * dst_dev, size_t dst_col_stride_bytes, void * stream) {
    if (!type_is_block(type)) return B200_ERR_TYPE;
    if (m < 0 || k < 0 || n < 0 || n > B200_MMVQ_MAX_N || k % type_qk(type) != 0) return B200_ERR_SHAPE;
    if (dst_col_stride_bytes % 4 != 0 || w_row_stride_bytes % 2 != 0) return B200_ERR_STRIDE;
    if (m == 0 || n == 0) return B200_OK;
    if (k == 0) return -(int) cudaMemset2DAsync(dst_dev, dst_col_stride_bytes, 0, (size_t) m * 4, (size_t) n, (cudaStream_t) stream);
    mmvq_args a;
    const size_t qbytes = ((size_t) n * (size_t) k + 255) / 256 * 256;
    a.w = (const uint8_t *) w_dev;
    a.yq = (const int8_t *) acts_scratch_dev;
    a.yd = (const float *) ((const uint8_t *) acts_scratch_dev + qbytes);
    a.dst = dst_dev;
    a.m = m; a.k = k; a.w_row_stride = (int64_t) w_row_stride_bytes; a.w_nb2 = 0; a.w_nb3 = 0;
    a.dst_col_stride = (int64_t) (dst_col_stride_bytes / 4); a.dst_nb2 = 0; a.dst_nb3 = 0;
    a.ne12 = 1; a.ne13 = 1; a.r2 = 1; a.r3 = 1; a.n = n; a.col_stride_n = 0; a.ids = nullptr; a.bad = nullptr;
    return mmvq_dispatch(type, a, (cudaStream_t) stream);
}

// ---------------------------------------------------------------- the MUL_MAT dispatcher
extern "C" size_t b200_mul_mat_workspace_size(int type_a, const int64_t ne_a[4], const int64_t ne_b[4]) {
    if (!type_is_block(type_a)) return 0;
    const int64_t k = ne_a[0], m = ne_a[1], n = ne_b[1];
    const int64_t cols = ne_b[1] * ne_b[2] * ne_b[3];
    if (k <= 0 || cols <= 0) return 256;
    if (n <= B200_MMVQ_MAX_N || !type_is_quant(type_a)) return b200_acts_scratch_bytes(k, cols) + 256;
    return b200_mul_mat_q_workspace_size(m, k, cols) + 256;
}

extern "C" int b200_mul_mat(int type_a, const void * a_dev, const int64_t ne_a[4], const size_t nb_a[4],
                            const float * b_dev, const int64_t ne_b[4], const size_t nb_b[4],
                            float * dst_dev, void * workspace_dev, size_t workspace_bytes, void * stream) {
    return b200_mul_mat_ex(type_a, a_dev, ne_a, nb_a, b_dev, ne_b, nb_b, dst_dev, workspace_dev, workspace_bytes, 0, stream);
}

extern "C" int b200_mul_mat_ex(int type_a, const void * a_dev, const int64_t ne_a[4], const size_t nb_a[4],
                               const float * b_dev, const int64_t ne_b[4], const size_t nb_b[4],
                               float * dst_dev, void * workspace_dev, size_t workspace_bytes, uint32_t flags, void * stream) {
    if (!type_is_block(type_a)) return B200_ERR_TYPE;
    const int64_t k = ne_a[0], m = ne_a[1], n = ne_b[1];
    for (int i = 0; i < 4; i++) if (ne_a[i] < 0 || ne_b[i] < 0) return B200_ERR_SHAPE;
    if (ne_b[0] != k || k % type_qk(type_a) != 0) return B200_ERR_SHAPE;
    if (ne_a[2] == 0 || ne_a[3] == 0) return (ne_b[2] == 0 || ne_b[3] == 0 || m == 0 || n == 0) ? B200_OK : B200_ERR_SHAPE;
    if (ne_b[2] % ne_a[2] != 0 || ne_b[3] % ne_a[3] != 0) return B200_ERR_SHAPE;
    if (nb_a[0] != (size_t) type_bytes(type_a) || nb_b[0] != sizeof(float)) return B200_ERR_STRIDE;
    if (nb_b[1] % 4 || nb_b[2] % 4 || nb_b[3] % 4) return B200_ERR_STRIDE;
    const int64_t ne12 = ne_b[2], ne13 = ne_b[3];
    if (m == 0 || n == 0 || ne12 == 0 || ne13 == 0) return B200_OK;
    cudaStream_t st = (cudaStream_t) stream;
    if (k == 0) return -(int) cudaMemsetAsync(dst_dev, 0, (size_t) (m * n * ne12 * ne13) * 4, st);
    if (workspace_bytes < b200_mul_mat_workspace_size(type_a, ne_a, ne_b) || !workspace_dev) return B200_ERR_WORKSPACE;
    uint8_t * ws = (uint8_t *) (((uintptr_t) workspace_dev + 255) & ~(uintptr_t) 255);

    if (n == 1 && ne12 == 1 && ne13 == 1 && type_is_quant(type_a)) {
        // decode fast path: activation quantization fused into the mat-vec, one launch
        const int rc = mmvq_fused_try(type_a, (const uint8_t *) a_dev, (int64_t) nb_a[1], m, k, b_dev, dst_dev,
                                      (flags & B200_MUL_MAT_CONST_WEIGHTS) != 0, st);
        if (rc != B200_STREAM_NOT_APPLICABLE) return rc;
    }
    if (n <= B200_MMVQ_MAX_N || !type_is_quant(type_a)) {
        // quantize every activation column once: scratch column index = (i13*ne12 + i12)*n + j
        // (Q4_1 / Q5_0 / Q5_1 / IQ4_NL have no tensor-core kernel: more than 8 columns run as groups of 8 on the general mat-vec)
        const int64_t cols = n * ne12 * ne13;
        const size_t qbytes = ((size_t) cols * (size_t) k + 255) / 256 * 256;
        int8_t * yq = (int8_t *) ws;
        float  * yd = (float *) (ws + qbytes);
        for (int64_t i13 = 0; i13 < ne13; i13++) for (int64_t i12 = 0; i12 < ne12; i12++) {
            const int64_t c0 = (i13 * ne12 + i12) * n;
            const float * src = (const float *) ((const uint8_t *) b_dev + i12 * nb_b[2] + i13 * nb_b[3]);
            int rc = b200_quantize_acts_planes(src, nb_b[1], k, n, yq + c0 * k, yd + c0 * (k / 32), st);
            if (rc) return rc;
        }
        mmvq_args a;
        a.w = (const uint8_t *) a_dev; a.yq = yq; a.yd = yd; a.dst = dst_dev;
        a.m = m; a.k = k; a.w_row_stride = (int64_t) nb_a[1]; a.w_nb2 = (int64_t) nb_a[2]; a.w_nb3 = (int64_t) nb_a[3];
        a.dst_col_stride = m; a.dst_nb2 = m * n; a.dst_nb3 = m * n * ne12;
        a.ne12 = ne12; a.ne13 = ne13; a.r2 = ne12 / ne_a[2]; a.r3 = ne13 / ne_a[3]; a.n = n; a.col_stride_n = 0; a.ids = nullptr; a.bad = nullptr;
        if (n <= B200_MMVQ_MAX_N) return mmvq_dispatch(type_a, a, st);
        // groups of 8 columns of every channel: a channel's columns sit n apart in the scratch, its dst m * n apart
        for (int64_t j0 = 0; j0 < n; j0 += B200_MMVQ_MAX_N) {
            mmvq_args g = a;
            g.n = n - j0 < B200_MMVQ_MAX_N ? n - j0 : B200_MMVQ_MAX_N;
            g.yq = yq + j0 * k; g.yd = yd + j0 * (k / 32); g.dst = dst_dev + j0 * m;
            g.col_stride_n = n;
            const int rc = mmvq_dispatch(type_a, g, st);
            if (rc) return rc;
        }
        return B200_OK;
    }
    // batched path, one 2-D problem per (i12, i13) channel
    for (int64_t i13 = 0; i13 < ne13; i13++) for (int64_t i12 = 0; i12 < ne12; i12++) {
        const uint8_t * w = (const uint8_t *) a_dev + (i12 / (ne12 / ne_a[2])) * nb_a[2] + (i13 / (ne13 / ne_a[3])) * nb_a[3];
        const float * x = (const float *) ((const uint8_t *) b_dev + i12 * nb_b[2] + i13 * nb_b[3]);
        float * d = dst_dev + (i13 * ne12 + i12) * m * n;
        int rc = b200_mul_mat_q(type_a, w, nb_a[1], m, k, x, nb_b[1], n, d, (size_t) m * 4, ws,
                                workspace_bytes - (size_t) (ws - (uint8_t *) workspace_dev), st);
        if (rc) return rc;
    }
    return B200_OK;
}

// ---------------------------------------------------------------- MUL_MAT_ID (mixture of experts)
// dst[m, e, t] = as[:, :, ids[e, t]] x b[:, e % ne_b[1], t]: every (slot, token) pair is a one-column mat-vec on the expert the
// ids tensor names.  The ids stay on the device (the reference copies them to the host and synchronizes for its
// batched path, src/ggml-cuda/ggml-cuda.cu:2092-2233; its decode path passes them to the kernel, mmvq.cu:163-166, as here).
extern "C" size_t b200_mul_mat_id_workspace_size(int type_a, const int64_t ne_a[4], const int64_t ne_b[4]) {
    if (!type_is_block(type_a) || ne_a[0] <= 0) return 256;
    return b200_acts_scratch_bytes(ne_a[0], ne_b[1] * ne_b[2]) + 512;
}
extern "C" int b200_mul_mat_id(int type_a, const void * as_dev, const int64_t ne_a[4], const size_t nb_a[4],
                               const float * b_dev, const int64_t ne_b[4], const size_t nb_b[4],
                               const int32_t * ids_dev, const int64_t ne_ids[2], const size_t nb_ids[2],
                               float * dst_dev, void * workspace_dev, size_t workspace_bytes, void * stream) {
    if (!type_is_block(type_a)) return B200_ERR_TYPE;
    const int64_t k = ne_a[0], m = ne_a[1], n_as = ne_a[2];
    const int64_t ne11 = ne_b[1], n_tok = ne_b[2], n_used = ne_ids[0];
    if (k < 0 || m < 0 || n_as < 1 || ne_a[3] != 1 || ne_b[3] != 1 || ne_b[0] != k || k % type_qk(type_a) != 0) return B200_ERR_SHAPE;
    if (ne_ids[1] != n_tok || ne11 < 1 || n_used < 1 || n_used % ne11 != 0 || n_used > 65535 || n_tok > 65535) return n_tok == 0 || n_used == 0 ? B200_OK : B200_ERR_SHAPE;
    if (nb_a[0] != (size_t) type_bytes(type_a) || nb_b[0] != sizeof(float) || nb_b[1] % 4 || nb_b[2] % 4 || nb_ids[0] % 4 || nb_ids[1] % 4) return B200_ERR_STRIDE;
    if (m == 0 || n_tok == 0) return B200_OK;
    cudaStream_t st = (cudaStream_t) stream;
    if (k == 0) return -(int) cudaMemsetAsync(dst_dev, 0, (size_t) (m * n_used * n_tok) * 4, st);
    if (!workspace_dev || workspace_bytes < b200_mul_mat_id_workspace_size(type_a, ne_a, ne_b)) return B200_ERR_WORKSPACE;
    uint8_t * ws = (uint8_t *) (((uintptr_t) workspace_dev + 255) & ~(uintptr_t) 255);
    const int64_t cols = ne11 * n_tok;
    const size_t qbytes = ((size_t) cols * (size_t) k + 255) / 256 * 256;
    int8_t * yq = (int8_t *) ws;
    float  * yd = (float *) (ws + qbytes);
    for (int64_t t = 0; t < n_tok; t++) {                      // scratch column index = t * ne11 + row of b
        if (nb_b[2] == (size_t) ne11 * nb_b[1] && t > 0) break;   // token slabs back to back: one call covered them all
        const float * src = (const float *) ((const uint8_t *) b_dev + t * nb_b[2]);
        const int64_t rows = nb_b[2] == (size_t) ne11 * nb_b[1] ? cols : ne11;
        const int rc = b200_quantize_acts_planes(src, nb_b[1], k, rows, yq + t * ne11 * k, yd + t * ne11 * (k / 32), st);
        if (rc) return rc;
    }
    mmvq_args a;
    a.w = (const uint8_t *) as_dev; a.yq = yq; a.yd = yd; a.dst = dst_dev;
    a.m = m; a.k = k; a.w_row_stride = (int64_t) nb_a[1]; a.w_nb2 = (int64_t) nb_a[2]; a.w_nb3 = 0;
    a.dst_col_stride = m; a.dst_nb2 = m; a.dst_nb3 = m * n_used;
    a.ne12 = n_used; a.ne13 = n_tok; a.r2 = 1; a.r3 = 1; a.n = 1; a.col_stride_n = 0;
    a.ids = (const uint8_t *) ids_dev; a.ids_nb0 = (int64_t) nb_ids[0]; a.ids_nb1 = (int64_t) nb_ids[1]; a.n_as = n_as; a.ne11 = ne11;
    a.bad = b200_error_word_dev();
    int rc;
    // decode (and small batches): the K-stationary TMA-staged kernel, one grid plane per (slot, token), expert looked up on the device
    // (the one-warp-per-row kernel below reached 25 % of the HBM peak on Mixtral-like experts)
    static const int moe_stream = env_int("GGML_B200_MOE_STREAM", 1);
    if (moe_stream && type_is_quant(type_a) && type_a != T_Q6_K && m >= 1) {          // Q6_K measured slower there (69 vs 52 us)
        stream_moe mo = { (const uint8_t *) ids_dev, (int64_t) nb_ids[0], (int64_t) nb_ids[1], (int64_t) nb_a[2], n_as, ne11, n_used, n_tok, a.bad };
        rc = mmvq_stream_try(type_a, (const uint8_t *) as_dev, (int64_t) nb_a[1], m, k, yq, yd, 1, dst_dev, m, st, &mo);
        if (rc == B200_OK) return rc;
        if (rc != B200_STREAM_NOT_APPLICABLE) return rc;
    }
    switch (type_a) {
        case T_Q4_0: rc = launch_generic<T_Q4_0>(a, st); break;
        case T_Q8_0: rc = launch_generic<T_Q8_0>(a, st); break;
        case T_Q4_K: rc = launch_generic<T_Q4_K>(a, st); break;
        case T_Q5_K: rc = launch_generic<T_Q5_K>(a, st); break;
        case T_Q4_1: rc = launch_generic<T_Q4_1>(a, st); break;
        case T_Q5_0: rc = launch_generic<T_Q5_0>(a, st); break;
        case T_Q5_1: rc = launch_generic<T_Q5_1>(a, st); break;
        case T_IQ4_NL: rc = launch_generic<T_IQ4_NL>(a, st); break;
        default:     rc = launch_generic<T_Q6_K>(a, st); break;
    }
    if (rc != B200_OK) return rc;
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

// ---------------------------------------------------------------- a list of MUL_MAT nodes
constexpr int LIST_LANES = 3;        // independent batched nodes of a list run on up to three streams (side_lanes below)
extern "C" size_t b200_mul_mat_batch_workspace_size(const b200_mul_mat_node * nodes, int64_t count) {
    // the largest single-node need + one q8 activation record per decode node (persistent launches) + the
    // completion counters of the (at most three, one per format family) launches of a flush
    size_t need = 256, acts = 0, batched = 0;
    int n_batched = 0;
    for (int64_t i = 0; i < count; i++) {
        const size_t s = b200_mul_mat_workspace_size(nodes[i].type_a, nodes[i].ne_a, nodes[i].ne_b);
        if (s > need) need = s;
        if (nodes[i].ne_b[1] > B200_MMVQ_MAX_N) { n_batched++; if (s > batched) batched = s; }      // may share the lanes (side_lanes)
        if (nodes[i].ne_b[1] >= 1 && nodes[i].ne_b[1] <= B200_MMVQ_MAX_N && nodes[i].ne_a[0] > 0) {
            const size_t per_col = (size_t) nodes[i].ne_b[1] * batch_acts_record_bytes(nodes[i].ne_a[0]);
            const size_t tile = nodes[i].ne_b[1] >= 2 ? cols_ws_bytes(nodes[i].ne_a[0], nodes[i].ne_b[1]) : 0;
            acts += per_col > tile ? per_col : tile;
        }
    }
    if (n_batched > 1 && LIST_LANES * (batched + 256) > need) need = LIST_LANES * (batched + 256);
    return need + acts + 3 * BATCH_DONE_BYTES + 256;
}

// What the pending launches read (x) and write (dst), indexed by address: a new node is checked against it for
// read-after-write, write-after-write and write-after-read.  Intervals are kept sorted by their start; none is longer
// than `maxlen`, so a query only walks the few neighbours that can overlap (a decode pass is ~230 nodes and this runs
// on the host every step outside a CUDA graph).
namespace {
struct hazard_index {
    struct iv { uintptr_t lo, hi; short owner, idx; };          // owner: 0..2 = format family builder, 3 = the MMA list kernel
    std::vector<iv> v;
    uintptr_t maxlen = 0;
    void clear() { v.clear(); maxlen = 0; }
    void add(uintptr_t lo, uintptr_t hi, int owner, int idx) {
        if (hi - lo > maxlen) maxlen = hi - lo;
        size_t a = 0, b = v.size();
        while (a < b) { const size_t mid = (a + b) / 2; if (v[mid].lo < lo) a = mid + 1; else b = mid; }
        // the same range again (q/k/v read one x): keep the latest owner index only
        if (a < v.size() && v[a].lo == lo && v[a].hi == hi && v[a].owner == owner) { if (idx > v[a].idx) v[a].idx = (short) idx; return; }
        v.insert(v.begin() + (ptrdiff_t) a, iv{ lo, hi, (short) owner, (short) idx });
    }
    // every interval overlapping [lo, hi): fn(owner, idx)
    template <typename F> void query(uintptr_t lo, uintptr_t hi, F fn) const {
        size_t a = 0, b = v.size();
        while (a < b) { const size_t mid = (a + b) / 2; if (v[mid].lo < hi) a = mid + 1; else b = mid; }   // first with lo >= hi
        for (size_t i = a; i-- > 0; ) {
            if (v[i].lo + maxlen <= lo) break;               // nothing further left can reach lo
            if (v[i].hi > lo) fn((int) v[i].owner, (int) v[i].idx);
        }
    }
};
struct batch_state {
    batch_builder * fam[3];
    cols_builder  * cols;
    hazard_index reads, writes;
    batch_state() { for (int f = 0; f < 3; f++) fam[f] = new batch_builder(f); cols = new cols_builder(); }
};
}

// plan != nullptr: dry run -- nothing is launched, no device is touched; plan[3 i .. 3 i + 2] = (launch group of node i,
// list index of the node it waits for inside that group or -1, route: 0 per-node call, 1 persistent list, 2 MMA list)
constexpr size_t PEER_CTRL_FAMILY = 8192;         // per format family (= launch of a flush): an epoch word, two sets of 448 node counters
                                                  // (one arrival per rank) and two sets of 448 local arrival counters (one per CTA)
constexpr size_t PEER_CTRL_BYTES = 32768;         // head of the symmetric region
static_assert(64 + 4 * BATCH_MAX_NODES * sizeof(unsigned) <= PEER_CTRL_FAMILY && 3 * PEER_CTRL_FAMILY <= PEER_CTRL_BYTES, "control block");

// Independent batched (n > 8) nodes of a list -- wq / wk / wv, gate / up of a prefill graph -- run side by side: the caller's
// stream plus two side streams forked and joined by events (capturable: a captured list becomes parallel graph branches),
// each lane with its own third of the workspace.  A 1024 x 4096 matrix at n = 512 fills 64 of 148 SMs for 34 us on its own.
namespace {
struct side_lanes {
    cudaStream_t s[LIST_LANES - 1] = {};
    cudaEvent_t  fork = nullptr, join[LIST_LANES - 1] = {};
    bool ok = false, tried = false;
};
side_lanes & get_side_lanes() {
    static thread_local side_lanes lanes[64];       // per host thread and device: a backend is driven by one thread at a time
    side_lanes & L = lanes[current_device_slot()];
    if (!L.tried) {
        L.tried = true;
        bool ok = cudaEventCreateWithFlags(&L.fork, cudaEventDisableTiming) == cudaSuccess;
        for (int i = 0; i < LIST_LANES - 1 && ok; i++)
            ok = cudaStreamCreateWithFlags(&L.s[i], cudaStreamNonBlocking) == cudaSuccess &&
                 cudaEventCreateWithFlags(&L.join[i], cudaEventDisableTiming) == cudaSuccess;
        if (!ok) cudaGetLastError();
        L.ok = ok;
    }
    return L;
}
}

static int run_list(const b200_mul_mat_node * nodes, int64_t count, void * workspace_dev, size_t workspace_bytes, void * stream,
                    int32_t * plan, const b200_peer_group * pg = nullptr, long long remote_delta = 0) {
    if (count < 0 || (count > 0 && !nodes)) return B200_ERR_SHAPE;
    const bool dry = plan != nullptr;
    // row split (b200_mul_mat_batch_peers): the local copy of the symmetric region and the distances to the others
    long long peer_delta[BATCH_MAX_PEERS];
    int n_peers = 0;
    uintptr_t sym_lo = 0, sym_hi = 0;
    if (pg) {
        if (pg->n_ranks < 1 || pg->n_ranks > BATCH_MAX_PEERS + 1 || pg->rank < 0 || pg->rank >= pg->n_ranks || pg->bytes < PEER_CTRL_BYTES) return B200_ERR_SHAPE;
        for (int r = 0; r < pg->n_ranks; r++) {
            if (!pg->base[r] || ((uintptr_t) pg->base[r] & 127)) return B200_ERR_STRIDE;
            if (r != pg->rank) peer_delta[n_peers++] = (long long) ((intptr_t) pg->base[r] - (intptr_t) pg->base[pg->rank]);
        }
        sym_lo = (uintptr_t) pg->base[pg->rank] + PEER_CTRL_BYTES; sym_hi = (uintptr_t) pg->base[pg->rank] + pg->bytes;
    }
    int group = 0;
    if (dry) { workspace_dev = (void *) (uintptr_t) 0x1000; workspace_bytes = (size_t) 1 << 40; }
    cudaStream_t st = (cudaStream_t) stream;
    const bool persistent = !b200_option(B200_OPT_NO_BATCH_KERNEL);
    const bool chain = !b200_option(B200_OPT_NO_CHAIN);
    // one builder (28 KB parameter block, reused) per format family = kernel instantiation, one for the MMA list
    // kernel (2..8 columns); per thread, because different backends may be driven from different threads
    static thread_local batch_state * S = nullptr;
    if (!S) S = new batch_state();
    batch_builder ** fam_bb = S->fam;
    cols_builder * cols_bb = S->cols;
    hazard_index & reads = S->reads, & writes = S->writes;
    static const int use_cols = env_int("GGML_B200_COLS_KERNEL", 1);
    static const int colmax = env_int("GGML_B200_BATCH_COLS", 8);
    ws_cursor cur;
    cur.reset((uint8_t *) workspace_dev, workspace_bytes);
    auto reset_family = [&](int f) {
        fam_bb[f]->reset();
        if (pg) fam_bb[f]->set_peers((unsigned *) ((uint8_t *) pg->base[pg->rank] + f * PEER_CTRL_FAMILY), peer_delta, n_peers);
        if (remote_delta) { fam_bb[f]->args.n_peers = 1; fam_bb[f]->args.peer_delta[0] = remote_delta; }    // one-way: no control block, no epochs
    };
    for (int f = 0; f < 3; f++) reset_family(f);
    cols_bb->reset();
    reads.clear(); writes.clear();
    const int grid = dry ? 148 : current_caps().sm_count;
    auto nothing_pending = [&]() { return fam_bb[0]->empty() && fam_bb[1]->empty() && fam_bb[2]->empty() && cols_bb->empty(); };
    auto flush = [&]() -> int {
        reads.clear(); writes.clear();
        int rc = B200_OK;
        if (dry) {
            if (!nothing_pending()) group++;
            for (int f = 0; f < 3; f++) reset_family(f);
            cols_bb->reset();
            cur.reset((uint8_t *) workspace_dev, workspace_bytes);
            return rc;
        }
        for (int f = 0; f < 3 && rc == B200_OK; f++) {
            batch_builder * bb = fam_bb[f];
            if (bb->empty()) continue;
            // one node (a launch of its own): small ones take the per-node decode kernel (same per-row
            // arithmetic, one launch instead of two); from ~4 MB of Q4_K/Q5_K weights the persistent pair is
            // faster even alone (4096x11008: 11.2 us against 17.4 us, profiles/r01m_single_probe.txt).
            // option batch_single = 1 / 0 forces the persistent / the per-node kernel.
            const int single_mode = b200_option(B200_OPT_BATCH_SINGLE);
            const batch_node nd = bb->args.nodes[0];
            const bool single_persistent = bb->args.last_scatter >= 0 ? true : (single_mode >= 0 ? single_mode != 0
                                         : (f == 0 && (int64_t) nd.m * nd.row_bytes >= (int64_t) 4 << 20));
            if (bb->args.count == 1 && !single_persistent) {
                rc = mmvq_fused_try(nd.type, nd.w, nd.row_stride, nd.m, nd.k, nd.x, nd.dst, bb->all_const, st);
                if (rc == B200_STREAM_NOT_APPLICABLE) rc = batch_launch(*bb, st);
            } else {
                rc = batch_launch(*bb, st);
            }
            if (rc == B200_STREAM_NOT_APPLICABLE && bb->args.last_scatter >= 0) rc = B200_ERR_SHAPE;   // no per-node path scatters
            if (rc == B200_STREAM_NOT_APPLICABLE) {
                // rows too long for two ring stages: the nodes of this builder one by one, in list order
                rc = B200_OK;
                for (int j = 0; j < bb->args.count && rc == B200_OK; j++) {
                    const b200_mul_mat_node & o = nodes[bb->orig[j]];
                    const batch_node & bn = bb->args.nodes[j];
                    const int64_t ne_b[4] = { o.ne_b[0], 1, 1, 1 };            // an entry is one column of its node
                    rc = b200_mul_mat_ex(o.type_a, o.a_dev, o.ne_a, o.nb_a, bn.x, ne_b, o.nb_b, bn.dst, workspace_dev, workspace_bytes, o.flags, stream);
                }
            }
            reset_family(f);
        }
        if (rc == B200_OK && !cols_bb->empty()) { rc = cols_launch(*cols_bb, st); cols_bb->reset(); }
        // stream order: the next launch's quantizer starts after these launches have finished with the workspace
        cur.reset((uint8_t *) workspace_dev, workspace_bytes);
        return rc;
    };
    // batched nodes in flight on the lanes (see side_lanes): their activation and output ranges
    struct inflight { uintptr_t x0, x1, d0, d1; };
    inflight infl[LIST_LANES];
    int n_inflight = 0;
    static const int use_lanes = env_int("GGML_B200_LIST_LANES", 1);
    auto join_lanes = [&]() -> int {
        if (n_inflight > 1) {
            side_lanes & L = get_side_lanes();
            for (int j = 1; j < n_inflight; j++) {
                cudaError_t e = cudaEventRecord(L.join[j - 1], L.s[j - 1]);
                if (e == cudaSuccess) e = cudaStreamWaitEvent(st, L.join[j - 1], 0);
                if (e != cudaSuccess) return -(int) e;
            }
        }
        n_inflight = 0;
        return B200_OK;
    };
    for (int64_t i = 0; i < count; i++) {
        const b200_mul_mat_node & nd = nodes[i];
        const int64_t k = nd.ne_a[0], m = nd.ne_a[1];
        // a few columns (decode at batch 2..8) ride the persistent launch as one entry per column, back to back:
        // the CTA streams its slab of W once from HBM and finds it in L2 for the other columns (a node's slab per
        // CTA is 64-200 KB), which beats the per-node kernel's two launches per MUL_MAT up to GGML_B200_BATCH_COLS
        const int64_t ncols = nd.ne_b[1];
        const bool flat = nd.ne_b[0] == k && nd.ne_a[2] == 1 && nd.ne_a[3] == 1 && nd.ne_b[2] == 1 && nd.ne_b[3] == 1 &&
                          nd.nb_a[0] == (size_t) type_bytes(nd.type_a) && nd.nb_b[0] == sizeof(float);
        const bool plain = persistent && flat && type_is_quant(nd.type_a) && ncols >= 1 && (ncols == 1 || ncols <= colmax) &&
                           (ncols == 1 || nd.nb_b[1] % 16 == 0) &&
                           batch_node_ok(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, nd.b_dev);
        const bool to_cols = persistent && use_cols && flat && ncols >= 2 &&
                             cols_node_ok(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, nd.b_dev, (int64_t) nd.nb_b[1], ncols);
        if (!to_cols && !plain) {
            if (remote_delta || (pg && (uintptr_t) nd.dst_dev + 4 > sym_lo - PEER_CTRL_BYTES && (uintptr_t) nd.dst_dev < sym_hi)) return B200_ERR_SHAPE;   // cannot scatter this shape
            int rc = flush();
            if (rc) return rc;
            if (dry) { plan[3 * i] = group++; plan[3 * i + 1] = -1; plan[3 * i + 2] = 0; continue; }
            // a batched node that neither reads nor overwrites what the nodes in flight on the other lanes touch joins them
            const int64_t cols_all = nd.ne_b[1] * nd.ne_b[2] * nd.ne_b[3];
            const uintptr_t x0 = (uintptr_t) nd.b_dev, x1 = x0 + (size_t) nd.nb_b[3] * (size_t) nd.ne_b[3];
            const uintptr_t d0 = (uintptr_t) nd.dst_dev, d1 = d0 + (size_t) m * (size_t) cols_all * 4;
            const size_t lane_ws = (workspace_bytes / LIST_LANES) & ~(size_t) 255;
            bool lane_ok = use_lanes && flat && ncols > B200_MMVQ_MAX_N && type_is_quant(nd.type_a) && !pg &&
                           b200_mul_mat_workspace_size(nd.type_a, nd.ne_a, nd.ne_b) <= lane_ws;
            if (lane_ok) {
                for (int j = 0; j < n_inflight && lane_ok; j++) {
                    const inflight & f = infl[j];
                    if ((x0 < f.d1 && f.d0 < x1) || (d0 < f.d1 && f.d0 < d1) || (d0 < f.x1 && f.x0 < d1)) lane_ok = false;
                }
            }
            if (!lane_ok || n_inflight == LIST_LANES) { rc = join_lanes(); if (rc) return rc; }
            if (lane_ok) {
                side_lanes & L = get_side_lanes();
                if (!L.ok) lane_ok = false;
                else {
                    const int lane = n_inflight;
                    cudaStream_t ls = st;
                    if (lane > 0) {
                        ls = L.s[lane - 1];
                        const cudaError_t e = cudaStreamWaitEvent(ls, L.fork, 0);
                        if (e != cudaSuccess) return -(int) e;
                    } else {
                        // lane 0 is the caller's stream: the fork point of the group is everything enqueued before this node
                        const cudaError_t e = cudaEventRecord(L.fork, st);
                        if (e != cudaSuccess) return -(int) e;
                    }
                    rc = b200_mul_mat_ex(nd.type_a, nd.a_dev, nd.ne_a, nd.nb_a, nd.b_dev, nd.ne_b, nd.nb_b, nd.dst_dev,
                                         (uint8_t *) workspace_dev + (size_t) lane * lane_ws, lane_ws, nd.flags, ls);
                    if (rc) return rc;
                    infl[n_inflight++] = inflight{ x0, x1, d0, d1 };
                    continue;
                }
            }
            rc = b200_mul_mat_ex(nd.type_a, nd.a_dev, nd.ne_a, nd.nb_a, nd.b_dev, nd.ne_b, nd.nb_b, nd.dst_dev,
                                 workspace_dev, workspace_bytes, nd.flags, stream);
            if (rc) return rc;
            continue;
        }
        { const int rc = join_lanes(); if (rc) return rc; }      // list nodes follow on the caller's stream
        const uintptr_t x0 = (uintptr_t) nd.b_dev, x1 = x0 + (size_t) (ncols - 1) * nd.nb_b[1] + (size_t) k * 4;
        const uintptr_t d0 = (uintptr_t) nd.dst_dev, d1 = d0 + (size_t) m * ncols * 4;
        const int owner = to_cols ? 3 : batch_family(nd.type_a);
        // row split: a dst inside the symmetric region is scattered to every rank -- only the persistent list kernel does that
        const bool scatter = (pg && d0 >= sym_lo && d1 <= sym_hi) || remote_delta != 0;
        if (pg && !scatter && d1 > sym_lo - PEER_CTRL_BYTES && d0 < sym_hi) return B200_ERR_STRIDE;     // straddles the region / its control block
        if (scatter && (to_cols || ncols != 1)) return B200_ERR_SHAPE;
        // hazards against what is pending: inside the same persistent launch they become a dependency (dep = the
        // latest such node; a CTA finishes nodes in order, so its counter covers the earlier ones), anywhere else
        // (another launch of this flush, the MMA list kernel, several columns, chaining switched off) they cut the list
        int dep = -1;
        bool cut = false, x_written = false;
        auto on_hit = [&](int o, int idx) { if (o != owner || owner == 3 || !chain || ncols != 1) cut = true; else if (idx > dep) dep = idx; };
        writes.query(x0, x1, [&](int o, int idx) { x_written = true; on_hit(o, idx); });      // read after write
        writes.query(d0, d1, on_hit);                                                          // write after write
        reads.query(d0, d1, on_hit);                                                           // write after read
        if (cut) {
            const int rc = flush();
            if (rc) return rc;
            dep = -1; x_written = false;
        }
        if (to_cols) {
            if (cols_bb->full()) { const int rc = flush(); if (rc) return rc; }
            bool ok = cols_bb->add(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, nd.b_dev, (int64_t) nd.nb_b[1], ncols, nd.dst_dev, cur);
            if (!ok && !nothing_pending()) {
                const int rc = flush();
                if (rc) return rc;
                ok = cols_bb->add(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, nd.b_dev, (int64_t) nd.nb_b[1], ncols, nd.dst_dev, cur);
            }
            if (ok) {
                reads.add(x0, x1, 3, cols_bb->count() - 1); writes.add(d0, d1, 3, cols_bb->count() - 1);
                if (dry) { plan[3 * i] = group; plan[3 * i + 1] = -1; plan[3 * i + 2] = 2; }
                continue;
            }
            if (!plain) {
                int rc = flush();
                if (rc) return rc;
                if (dry) { plan[3 * i] = group++; plan[3 * i + 1] = -1; plan[3 * i + 2] = 0; continue; }
                rc = b200_mul_mat_ex(nd.type_a, nd.a_dev, nd.ne_a, nd.nb_a, nd.b_dev, nd.ne_b, nd.nb_b, nd.dst_dev,
                                     workspace_dev, workspace_bytes, nd.flags, stream);
                if (rc) return rc;
                continue;
            }
        }
        batch_builder * bb = fam_bb[batch_family(nd.type_a)];
        const bool cw = (nd.flags & B200_MUL_MAT_CONST_WEIGHTS) != 0;
        int64_t c = 0;
        for (; c < ncols; c++) {
            const float * xc = (const float *) ((const uint8_t *) nd.b_dev + c * nd.nb_b[1]);
            float * dc = nd.dst_dev + c * m;
            if (bb->full()) {
                const int rc = flush();
                if (rc) return rc;
                dep = -1; x_written = false;
            }
#ifdef B200_TUNING
            // timing experiments only (wrong results): 1 = wait but take pre-quantized x, 2 = quantize in the launch but do not wait
            static const int chain_dbg = env_int("GGML_B200_CHAIN_DBG", 0);
            if (chain_dbg == 1) x_written = false;
            if (chain_dbg == 2 && dep >= 0) { dep = -2; }
#endif
            bool ok = bb->add(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, xc, dc, cw, grid, cur, dep < 0 ? -1 : dep, x_written && dep != -1, (int) i, scatter);
            if (!ok && !nothing_pending()) {                   // workspace full: launch what is pending, start over
                const int rc = flush();
                if (rc) return rc;
                dep = -1; x_written = false;
                ok = bb->add(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, xc, dc, cw, grid, cur, -1, false, (int) i, scatter);
            }
            if (!ok) break;
        }
        if (c == ncols) {
            const int idx = bb->args.count - 1;
            reads.add(x0, x1, owner, idx); writes.add(d0, d1, owner, idx);
            if (dry) {
                const int dp = bb->args.nodes[idx].dep;
                plan[3 * i] = group; plan[3 * i + 1] = dp >= 0 ? bb->orig[dp] : -1; plan[3 * i + 2] = 1;
            }
            continue;
        }
        // no room even for one record (not reachable with a workspace of b200_mul_mat_batch_workspace_size): finish
        // what is pending, then the per-node path reports the workspace error or computes the node
        int rc = flush();
        if (rc) return rc;
        if (dry) { plan[3 * i] = group++; plan[3 * i + 1] = -1; plan[3 * i + 2] = 0; continue; }
        rc = b200_mul_mat_ex(nd.type_a, nd.a_dev, nd.ne_a, nd.nb_a, nd.b_dev, nd.ne_b, nd.nb_b, nd.dst_dev,
                             workspace_dev, workspace_bytes, nd.flags, stream);
        if (rc) return rc;
    }
    { const int rc = join_lanes(); if (rc) return rc; }
    return flush();
}

extern "C" int b200_mul_mat_batch(const b200_mul_mat_node * nodes, int64_t count,
                                  void * workspace_dev, size_t workspace_bytes, void * stream) {
    return run_list(nodes, count, workspace_dev, workspace_bytes, stream, nullptr);
}

// Row split: the same list on every rank's row blocks, the gather fused into the launch (mmvq_batch.cuh).
extern "C" size_t b200_peer_ctrl_bytes(void) { return PEER_CTRL_BYTES; }
extern "C" int b200_mul_mat_batch_peers(const b200_mul_mat_node * nodes, int64_t count, const b200_peer_group * peers,
                                        void * workspace_dev, size_t workspace_bytes, void * stream) {
    if (!peers) return B200_ERR_SHAPE;
    return run_list(nodes, count, workspace_dev, workspace_bytes, stream, nullptr, peers);
}

// One-way variant for a single process driving several devices (the ggml split buffer): every node's finished rows are
// ALSO stored at dst_dev + remote_delta_bytes, an address in another device's memory (peer access enabled by the caller).
// No cross-device protocol: the remote rows are complete when this call's work on `stream` is (the caller orders the
// reader with an event).  One-column decode nodes only (B200_ERR_SHAPE otherwise).
extern "C" int b200_mul_mat_batch_remote(const b200_mul_mat_node * nodes, int64_t count, int64_t remote_delta_bytes,
                                         void * workspace_dev, size_t workspace_bytes, void * stream) {
    if (remote_delta_bytes == 0 || remote_delta_bytes % 4 != 0) return B200_ERR_STRIDE;
    return run_list(nodes, count, workspace_dev, workspace_bytes, stream, nullptr, nullptr, (long long) remote_delta_bytes);
}

// How b200_mul_mat_batch would issue the list (host arithmetic only; no device needed).
extern "C" int b200_mul_mat_batch_plan(const b200_mul_mat_node * nodes, int64_t count, int32_t * plan) {
    if (!plan) return B200_ERR_SHAPE;
    return run_list(nodes, count, nullptr, 0, nullptr, plan);
}

// ---------------------------------------------------------------- the list kernels' quantizers on their own
// (tests compare their output bit for bit with quantize_row_q8_1_ref; the layouts are documented in the header)
extern "C" size_t b200_acts_record_bytes(int64_t k) { return k > 0 ? batch_acts_record_bytes(k) : 0; }
extern "C" size_t b200_acts_columns_bytes(int64_t k, int64_t n) { return k > 0 ? cols_ws_bytes(k, n) : 0; }

extern "C" int b200_quantize_acts_record(const float * x_dev, int64_t k, void * record_dev, void * stream) {
    if (k < 256 || k % 256 != 0 || k > 32768) return B200_ERR_SHAPE;
    if (((uintptr_t) x_dev & 15) || ((uintptr_t) record_dev & 127)) return B200_ERR_STRIDE;
    static thread_local batch_args * a = nullptr;
    if (!a) a = new batch_args();
    a->count = 1; a->kmax = (int) k; a->done = nullptr;
    batch_node & nd = a->nodes[0];
    memset(&nd, 0, sizeof(nd));
    nd.x = x_dev; nd.acts = (const uint8_t *) record_dev; nd.k = (int) k; nd.qflags = BATCH_Q_PRE | BATCH_Q_LOAD; nd.dep = -1;
    const dim3 grid((unsigned) ceil_div(k / 128, 8), 1);
    k_quantize_batch<<<grid, 256, 0, (cudaStream_t) stream>>>(*a);
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

extern "C" int b200_quantize_acts_columns(const float * x_dev, size_t x_row_stride_bytes, int64_t k, int64_t n, void * columns_dev, void * stream) {
    if (k < 256 || k % 256 != 0 || n < 1 || n > COLS_N) return B200_ERR_SHAPE;
    if (((uintptr_t) x_dev & 3) || (x_row_stride_bytes & 3) || ((uintptr_t) columns_dev & 127)) return B200_ERR_STRIDE;
    static thread_local cols_args * a = nullptr;
    if (!a) a = new cols_args();
    a->count = 1; a->units = 0; a->kmax = (int) k; a->nq = 1; a->qlist[0] = 0;
    cols_node & nd = a->nodes[0];
    memset(&nd, 0, sizeof(nd));
    nd.x = x_dev; nd.k = (int) k; nd.n = (int) n; nd.x_stride = (int) (x_row_stride_bytes / 4); nd.quantize = 1;
    nd.yq = (int8_t *) columns_dev;
    nd.ym = (uint8_t *) columns_dev + (((size_t) COLS_N * k + 127) & ~(size_t) 127);
    const dim3 grid((unsigned) ((k / 256 + 7) / 8), 1, COLS_N);
    k_quantize_cols<<<grid, 256, 0, (cudaStream_t) stream>>>(*a);
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

#ifdef B200_TUNING
// tuning builds only: every fused decode launch writes 8 globaltimer stamps per CTA into buf
extern "C" __attribute__((visibility("default"))) void b200_debug_timestamps(void * buf) { g_fused_dbg = (unsigned long long *) buf; }
extern "C" __attribute__((visibility("default"))) void b200_debug_batch_stamps(void * buf) { g_batch_dbg = (unsigned long long *) buf; }
#endif
