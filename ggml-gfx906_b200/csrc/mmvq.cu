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
    int64_t n;              // columns per channel
};

template <int T, int NCOLS>
__global__ void __launch_bounds__(128) k_mmvq_generic(const mmvq_args a) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row = (int64_t) blockIdx.x * 4 + warp;
    if (row >= a.m) return;
    const int64_t i12 = blockIdx.y, i13 = blockIdx.z;
    const uint8_t * wrow = a.w + row * a.w_row_stride + (i12 / a.r2) * a.w_nb2 + (i13 / a.r3) * a.w_nb3;
    const int64_t col0 = (i13 * a.ne12 + i12) * a.n;          // first column of this channel in the scratch
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
    // fast path: plain 2-D matrix, rows contiguous
    if (a.ne12 == 1 && a.ne13 == 1) {
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
    if (!type_is_quant(type)) return B200_ERR_TYPE;
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
    a.ne12 = 1; a.ne13 = 1; a.r2 = 1; a.r3 = 1; a.n = n;
    return mmvq_dispatch(type, a, (cudaStream_t) stream);
}

// ---------------------------------------------------------------- the MUL_MAT dispatcher
extern "C" size_t b200_mul_mat_workspace_size(int type_a, const int64_t ne_a[4], const int64_t ne_b[4]) {
    if (!type_is_quant(type_a)) return 0;
    const int64_t k = ne_a[0], m = ne_a[1], n = ne_b[1];
    const int64_t cols = ne_b[1] * ne_b[2] * ne_b[3];
    if (k <= 0 || cols <= 0) return 256;
    if (n <= B200_MMVQ_MAX_N) return b200_acts_scratch_bytes(k, cols) + 256;
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
    if (!type_is_quant(type_a)) return B200_ERR_TYPE;
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

    if (n == 1 && ne12 == 1 && ne13 == 1) {
        // decode fast path: activation quantization fused into the mat-vec, one launch
        const int rc = mmvq_fused_try(type_a, (const uint8_t *) a_dev, (int64_t) nb_a[1], m, k, b_dev, dst_dev,
                                      (flags & B200_MUL_MAT_CONST_WEIGHTS) != 0, st);
        if (rc != B200_STREAM_NOT_APPLICABLE) return rc;
    }
    if (n <= B200_MMVQ_MAX_N) {
        // quantize every activation column once: scratch column index = (i13*ne12 + i12)*n + j
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
        a.ne12 = ne12; a.ne13 = ne13; a.r2 = ne12 / ne_a[2]; a.r3 = ne13 / ne_a[3]; a.n = n;
        return mmvq_dispatch(type_a, a, st);
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

// ---------------------------------------------------------------- a list of MUL_MAT nodes
extern "C" size_t b200_mul_mat_batch_workspace_size(const b200_mul_mat_node * nodes, int64_t count) {
    // the largest single-node need + one q8 activation record per decode node (persistent launches)
    size_t need = 256, acts = 0;
    for (int64_t i = 0; i < count; i++) {
        const size_t s = b200_mul_mat_workspace_size(nodes[i].type_a, nodes[i].ne_a, nodes[i].ne_b);
        if (s > need) need = s;
        if (nodes[i].ne_b[1] >= 1 && nodes[i].ne_b[1] <= B200_MMVQ_MAX_N && nodes[i].ne_a[0] > 0) {
            const size_t per_col = (size_t) nodes[i].ne_b[1] * batch_acts_record_bytes(nodes[i].ne_a[0]);
            const size_t tile = nodes[i].ne_b[1] >= 2 ? cols_ws_bytes(nodes[i].ne_a[0], nodes[i].ne_b[1]) : 0;
            acts += per_col > tile ? per_col : tile;
        }
    }
    return need + acts + 256;
}

// Pointer ranges a pending persistent launch reads (x) and writes (dst): a node joins the launch only if
// it neither reads nor writes what an earlier node of the launch writes, nor writes what one reads.
namespace {
struct range_set {
    struct iv { uintptr_t lo, hi; };
    std::vector<iv> v;
    uintptr_t lo = ~(uintptr_t) 0, hi = 0;
    void clear() { v.clear(); lo = ~(uintptr_t) 0; hi = 0; }
    void add(uintptr_t a, uintptr_t b) { v.push_back({a, b}); if (a < lo) lo = a; if (b > hi) hi = b; }
    bool hits(uintptr_t a, uintptr_t b) const {
        if (a >= hi || b <= lo) return false;                 // outside the hull: the common case
        for (const iv & i : v) if (a < i.hi && i.lo < b) return true;
        return false;
    }
};
}

extern "C" int b200_mul_mat_batch(const b200_mul_mat_node * nodes, int64_t count,
                                  void * workspace_dev, size_t workspace_bytes, void * stream) {
    if (count < 0 || (count > 0 && !nodes)) return B200_ERR_SHAPE;
    cudaStream_t st = (cudaStream_t) stream;
    const bool persistent = !env_int("GGML_B200_NO_BATCH_KERNEL", 0);
    // one builder (28 KB parameter block, reused) per format family = kernel instantiation; the nodes of a
    // pending set are independent, so sorting them into per-family launches keeps the semantics
    static thread_local batch_builder * fam_bb[3] = { nullptr, nullptr, nullptr };
    static thread_local range_set reads, writes;
    static thread_local cols_builder * cols_bb = nullptr;       // nodes with 2..8 columns (Q4_K / Q5_K): the MMA list kernel
    if (!cols_bb) cols_bb = new cols_builder();
    cols_bb->reset();
    static const int use_cols = env_int("GGML_B200_COLS_KERNEL", 1);
    if (!fam_bb[0]) { for (int f = 0; f < 3; f++) fam_bb[f] = new batch_builder(f); }
    ws_cursor cur;
    cur.reset((uint8_t *) workspace_dev, workspace_bytes);
    for (int f = 0; f < 3; f++) fam_bb[f]->reset();
    reads.clear(); writes.clear();
    const int grid = current_caps().sm_count;
    auto flush = [&]() -> int {
        reads.clear(); writes.clear();
        int rc = B200_OK;
        for (int f = 0; f < 3 && rc == B200_OK; f++) {
            batch_builder * bb = fam_bb[f];
            if (bb->empty()) continue;
            // one node (a dependent chain's link): small ones take the per-node decode kernel (same per-row
            // arithmetic, one launch instead of two); from ~4 MB of Q4_K/Q5_K weights the persistent pair is
            // faster even alone (4096x11008: 11.2 us against 17.4 us, profiles/r01m_single_probe.txt).
            // GGML_B200_BATCH_SINGLE=1 / 0 forces the persistent / the per-node kernel.
            const int single_mode = env_int("GGML_B200_BATCH_SINGLE", -1);
            const batch_node nd = bb->args.nodes[0];
            const bool single_persistent = single_mode >= 0 ? single_mode != 0
                                         : (f == 0 && (int64_t) nd.m * nd.row_bytes >= (int64_t) 4 << 20);
            if (bb->args.count == 1 && !single_persistent) {
                rc = mmvq_fused_try(nd.type, nd.w, nd.row_stride, nd.m, nd.k, nd.x, nd.dst, bb->all_const, st);
                if (rc == B200_STREAM_NOT_APPLICABLE) rc = batch_launch(*bb, st);
            } else {
                rc = batch_launch(*bb, st);
            }
            bb->reset();
        }
        if (rc == B200_OK && !cols_bb->empty()) { rc = cols_launch(*cols_bb, st); cols_bb->reset(); }
        // stream order: the next launch's quantizer starts after these launches have finished with the workspace
        cur.reset((uint8_t *) workspace_dev, workspace_bytes);
        return rc;
    };
    for (int64_t i = 0; i < count; i++) {
        const b200_mul_mat_node & nd = nodes[i];
        const int64_t k = nd.ne_a[0], m = nd.ne_a[1];
        // a few columns (decode at batch 2..8) ride the persistent launch as one entry per column, back to back:
        // the CTA streams its slab of W once from HBM and finds it in L2 for the other columns (a node's slab per
        // CTA is 64-200 KB), which beats the per-node kernel's two launches per MUL_MAT up to GGML_B200_BATCH_COLS
        static const int colmax = env_int("GGML_B200_BATCH_COLS", 8);
        const int64_t ncols = nd.ne_b[1];
        const bool plain = persistent && type_is_quant(nd.type_a) && nd.ne_b[0] == k && ncols >= 1 && (ncols == 1 || ncols <= colmax) &&
                           (ncols == 1 || nd.nb_b[1] % 16 == 0) &&
                           nd.ne_a[2] == 1 && nd.ne_a[3] == 1 && nd.ne_b[2] == 1 && nd.ne_b[3] == 1 &&
                           nd.nb_a[0] == (size_t) type_bytes(nd.type_a) && nd.nb_b[0] == sizeof(float) &&
                           batch_node_ok(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, nd.b_dev);
        if (persistent && use_cols && ncols >= 2 && nd.ne_b[0] == k && nd.ne_a[2] == 1 && nd.ne_a[3] == 1 && nd.ne_b[2] == 1 && nd.ne_b[3] == 1 &&
            nd.nb_a[0] == (size_t) type_bytes(nd.type_a) && nd.nb_b[0] == sizeof(float) &&
            cols_node_ok(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, nd.b_dev, (int64_t) nd.nb_b[1], ncols)) {
            const uintptr_t x0 = (uintptr_t) nd.b_dev, x1 = x0 + (size_t) (ncols - 1) * nd.nb_b[1] + (size_t) k * 4;
            const uintptr_t d0 = (uintptr_t) nd.dst_dev, d1 = d0 + (size_t) m * ncols * 4;
            if (cols_bb->full() || writes.hits(x0, x1) || writes.hits(d0, d1) || reads.hits(d0, d1)) {
                const int rc = flush();
                if (rc) return rc;
            }
            bool ok = cols_bb->add(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, nd.b_dev, (int64_t) nd.nb_b[1], ncols, nd.dst_dev, cur);
            if (!ok && !(fam_bb[0]->empty() && fam_bb[1]->empty() && fam_bb[2]->empty() && cols_bb->empty())) {
                const int rc = flush();
                if (rc) return rc;
                ok = cols_bb->add(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, nd.b_dev, (int64_t) nd.nb_b[1], ncols, nd.dst_dev, cur);
            }
            if (ok) { reads.add(x0, x1); writes.add(d0, d1); continue; }
        }
        if (plain) {
            const uintptr_t x0 = (uintptr_t) nd.b_dev, x1 = x0 + (size_t) (ncols - 1) * nd.nb_b[1] + (size_t) k * 4;
            const uintptr_t d0 = (uintptr_t) nd.dst_dev, d1 = d0 + (size_t) m * ncols * 4;
            batch_builder * bb = fam_bb[batch_family(nd.type_a)];
            if (writes.hits(x0, x1) || writes.hits(d0, d1) || reads.hits(d0, d1)) {
                const int rc = flush();
                if (rc) return rc;
            }
            const bool cw = (nd.flags & B200_MUL_MAT_CONST_WEIGHTS) != 0;
            int64_t c = 0;
            for (; c < ncols; c++) {
                const float * xc = (const float *) ((const uint8_t *) nd.b_dev + c * nd.nb_b[1]);
                float * dc = nd.dst_dev + c * m;
                if (bb->full()) {
                    const int rc = flush();
                    if (rc) return rc;
                }
                bool ok = bb->add(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, xc, dc, cw, grid, cur);
                if (!ok && !(fam_bb[0]->empty() && fam_bb[1]->empty() && fam_bb[2]->empty() && cols_bb->empty())) {  // workspace full: launch what is pending, start over
                    const int rc = flush();
                    if (rc) return rc;
                    ok = bb->add(nd.type_a, (const uint8_t *) nd.a_dev, (int64_t) nd.nb_a[1], m, k, xc, dc, cw, grid, cur);
                }
                if (!ok) break;
            }
            if (c == ncols) { reads.add(x0, x1); writes.add(d0, d1); continue; }
            if (c > 0) {                              // some columns are pending and the rest has no room: not reachable with a
                const int rc = flush();               // workspace of b200_mul_mat_batch_workspace_size; finish them, then recompute the node
                if (rc) return rc;
            }
            // no room even for one record: the per-node path below reports the workspace error
        }
        int rc = flush();
        if (rc) return rc;
        rc = b200_mul_mat_ex(nd.type_a, nd.a_dev, nd.ne_a, nd.nb_a, nd.b_dev, nd.ne_b, nd.nb_b, nd.dst_dev,
                             workspace_dev, workspace_bytes, nd.flags, stream);
        if (rc) return rc;
    }
    return flush();
}

// tuning aid used by tools/: every fused decode launch writes 8 globaltimer stamps per CTA into buf
extern "C" __attribute__((visibility("default"))) void b200_debug_timestamps(void * buf) { g_fused_dbg = (unsigned long long *) buf; }
