/*
 * ggml_b200.h — the C ABI of libggml_b200_kernels.so: hand-written sm_100a CUDA for ggml's
 * quantized hot path (Q4_0 / Q8_0 / Q4_K / Q5_K / Q6_K dequantization, q8_1 activation
 * quantization, GET_ROWS, MUL_MAT).  Plain pointers and sizes only; every pointer named *_dev is a
 * device pointer on the current CUDA device; `stream` is a cudaStream_t passed as void* (NULL =
 * default stream).  All calls are asynchronous on `stream` unless stated otherwise.
 *
 * Return value: 0 on success, B200_ERR_* (> 0) for argument errors, -(cudaError_t) for CUDA errors.
 * There is no CPU fallback anywhere behind this ABI: without a usable device the calls fail.
 *
 * Each entry point names the reference interface it replaces (paths relative to the reference
 * tree skyne98/ggml-gfx906).  Tensor conventions are ggml's (include/ggml.h:614-646): ne[] are
 * extents in elements, nb[] are strides in BYTES, dimension 0 is the contiguous one.
 */
#ifndef GGML_B200_H
#define GGML_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define B200_API __declspec(dllexport)
#else
#define B200_API __attribute__((visibility("default")))
#endif

/* type ids = the numeric values of enum ggml_type (include/ggml.h:376-410) */
enum b200_type {
    B200_TYPE_F32  = 0,
    B200_TYPE_F16  = 1,
    B200_TYPE_Q4_0 = 2,
    B200_TYPE_Q4_1 = 3,   /* Q4_1, Q5_0, Q5_1, IQ4_NL: dequantize / GET_ROWS / CPY and the general mat-vec (MUL_MAT, MUL_MAT_ID) only */
    B200_TYPE_Q5_0 = 6,
    B200_TYPE_Q5_1 = 7,
    B200_TYPE_Q8_0 = 8,
    B200_TYPE_Q8_1 = 9,
    B200_TYPE_Q4_K = 12,
    B200_TYPE_Q5_K = 13,
    B200_TYPE_Q6_K = 14,
    B200_TYPE_IQ4_NL = 20,
    B200_TYPE_I32  = 26,
    B200_TYPE_BF16 = 30,
};

enum b200_error {
    B200_OK              = 0,
    B200_ERR_TYPE        = 1, /* type not one of the five weight formats / unsupported output type */
    B200_ERR_SHAPE       = 2, /* k not a multiple of the block size, negative extents, bad broadcast */
    B200_ERR_STRIDE      = 3, /* dimension 0 not contiguous, or misaligned strides */
    B200_ERR_WORKSPACE   = 4, /* workspace too small / NULL */
    B200_ERR_NO_DEVICE   = 5, /* no sm_100 device visible */
    B200_ERR_INDEX       = 6, /* a GET_ROWS index outside the table (reported by b200_device_error_flags) */
    B200_ERR_TOO_LARGE   = 7, /* the problem needs more than the fixed static scratch (b200_mul_mat_vec_static) */
    B200_ERR_FILE        = 8, /* a GGUF file could not be opened, parsed or read (text: b200_gguf_last_error) */
    B200_NOT_APPLICABLE  = 1000, /* internal: the fast path declined; never returned by the entry points below */
};

/* ---- library / device ---------------------------------------------------------------------
 * replaces ggml_backend_cuda_get_device_count / _get_device_description / _get_device_memory
 * (include/ggml-cuda.h:36-40) and the arch probe in ggml_cuda_init (src/ggml-cuda/ggml-cuda.cu:196-306). */
B200_API int          b200_abi_version(void);
B200_API const char * b200_error_string(int code);
B200_API int          b200_device_count(void);                 /* devices with compute capability 10.x */
B200_API int          b200_device_info(int device, char * name, size_t name_len,
                                       size_t * free_bytes, size_t * total_bytes, int * sm_count, int * cc);

/* ---- run-time path switches ----------------------------------------------------------------
 * Which kernel serves a call is chosen by shape; tests and benches force the other candidates to
 * compare them on the same inputs.  Each option starts from its environment variable (read once,
 * at first use); b200_set_option returns the previous value.  No launch path reads the environment.
 * (replaces the GGML_CUDA_FORCE_* compile-time switches, src/ggml-cuda/ggml-cuda.cu:2029-2060.) */
enum b200_option {
    B200_OPT_MMVQ_GENERIC        = 0, /* GGML_B200_MMVQ_GENERIC=1: one-warp-per-row kernel for every n <= 8 call */
    B200_OPT_MMVQ_NO_FUSED       = 1, /* GGML_B200_MMVQ_NO_FUSED=1: n = 1 without the fused quantize + mat-vec kernel */
    B200_OPT_BATCH_SINGLE        = 2, /* GGML_B200_BATCH_SINGLE=1/0: a one-node list takes the persistent / the per-node kernel (-1: by size) */
    B200_OPT_NO_BATCH_KERNEL     = 3, /* GGML_B200_NO_BATCH_KERNEL=1: lists run as one launch per node */
    B200_OPT_MMQ_BN              = 4, /* GGML_B200_MMQ_BN=32|64|128|256|512: tile width of the batched kernel */
    B200_OPT_MMQ_FORCE_UNALIGNED = 5, /* GGML_B200_MMQ_FORCE_UNALIGNED=1: per-row raw copies instead of TMA slab tiles */
    B200_OPT_NO_CHAIN            = 6, /* GGML_B200_NO_CHAIN=1: a list is cut at every read-after-write instead of chaining inside one launch */
    B200_OPT_MMQ_ROUTE           = 7, /* GGML_B200_MMQ_ROUTE=1|2: batched MUL_MAT through the bf16 / the q8_1 int8 tensor-core kernel (0: by shape) */
    B200_OPT_COUNT
};
B200_API int b200_set_option(int option, int value);

/* Sticky error bits set by kernels that met bad data (bit 0: B200_ERR_INDEX); valid once the stream that ran
 * the kernel is synchronized.  replaces the GGML_ASSERT on the row index in ggml_compute_forward_get_rows_q
 * (src/ggml-cpu/ops.cpp:5147).  clear != 0 resets the word. */
B200_API unsigned b200_device_error_flags(int clear);

/* ---- format table -------------------------------------------------------------------------
 * replaces ggml_blck_size / ggml_type_size / ggml_row_size for the hot-path types (include/ggml.h:716-721) */
B200_API int    b200_block_elems(int type);
B200_API size_t b200_block_bytes(int type);
B200_API size_t b200_row_bytes(int type, int64_t k);

/* ---- dequantize_row_* ----------------------------------------------------------------------
 * replaces to_fp32_cuda_t / to_fp16_cuda_t from ggml_get_to_fp32_cuda / ggml_get_to_fp16_cuda
 * (src/ggml-cuda/convert.cu:661-764); values bit-exact with dequantize_row_<type>
 * (src/ggml-quants.c:307-325, 401-415, 1352-1374, 1554-1579, 1762-1791) and, for f16 output,
 * with the RN-even narrowing of those f32 values.
 * x_dev: n_elems/QK contiguous blocks.  out_type: B200_TYPE_F32, B200_TYPE_F16 or B200_TYPE_BF16 (RN-even narrowing:
 * ggml_get_to_bf16_cuda, src/ggml-cuda/convert.cu:650-659).  Also Q4_1 / Q5_0 / Q5_1 / IQ4_NL (src/ggml-quants.c:327-399, 2512-2528). */
B200_API int b200_dequantize_row(int type, const void * x_dev, void * y_dev, int64_t n_elems,
                                 int out_type, void * stream);

/* Strided 4-D form: CPY / DUP / CONT of a quantized tensor into an f32/f16 tensor.
 * replaces ggml_cuda_cpy for q -> f32 (src/ggml-cuda/cpy.cu:53-64, 94-121, 185-196) and the _nc_
 * converters (src/ggml-cuda/convert.cu:766-827).  src rows (dim 0) must be contiguous blocks;
 * dst element (i0,i1,i2,i3) is written at byte offset i0*nb_dst[0]+..+i3*nb_dst[3]; the logical
 * element order of src and dst is the same (ggml CPY semantics, shapes may differ as long as the
 * element counts match and ne_dst[0] % 8 == 0 or ne_src == ne_dst). */
B200_API int b200_dequantize_4d(int type, const void * x_dev, const int64_t ne_src[4], const size_t nb_src[4],
                                void * y_dev, int out_type, const int64_t ne_dst[4], const size_t nb_dst[4],
                                void * stream);

/* ---- quantize_row_q8_1 ---------------------------------------------------------------------
 * replaces quantize_row_q8_1_cuda (src/ggml-cuda/quantize.cu:145-157) with the value contract of
 * quantize_row_q8_1_ref (src/ggml-quants.c:225-258): bit-exact block_q8_1 output.
 * x_dev: n rows of k floats, row stride x_row_stride_bytes; y_dev: n * k/32 block_q8_1, rows packed. */
B200_API int b200_quantize_row_q8_1(const float * x_dev, size_t x_row_stride_bytes, int64_t k, int64_t n,
                                    void * y_dev, void * stream);

/* ---- quantize rows: f32 -> Q8_0 / Q4_0 (CPY f32 -> quant) ------------------------------------
 * replaces ggml_cuda_cpy for f32 -> q8_0 / q4_0 (src/ggml-cuda/cpy.cu:66-92, 174-214) with the value contract of the CPU
 * definitions quantize_row_q8_0_ref / quantize_row_q4_0_ref (src/ggml-quants.c:199-222, 36-71): bit-exact blocks.
 * x_dev: n rows of k floats (k % 32 == 0); y_dev: n rows of k / 32 blocks, rows y_row_stride_bytes apart. */
B200_API int b200_quantize_rows(int type, const float * x_dev, size_t x_row_stride_bytes, int64_t k, int64_t n,
                                void * y_dev, size_t y_row_stride_bytes, void * stream);
/* Strided 4-D form (ggml CPY between tensors of one shape, e.g. a permuted f32 source into a K/V-cache view): ne = the shape
 * in elements, nb_x / nb_y = ggml byte strides; dim 0 dense on both sides (nb_x[0] == 4, nb_y[0] == block bytes). */
B200_API int b200_quantize_rows_4d(int type, const float * x_dev, const int64_t ne[4], const size_t nb_x[4],
                                   void * y_dev, const size_t nb_y[4], void * stream);

/* ---- GET_ROWS ------------------------------------------------------------------------------
 * replaces ggml_cuda_op_get_rows / get_rows_cuda_q (src/ggml-cuda/getrows.cu:93-124, 236-253) and
 * extends it to the K-quants, which the reference serves only on the CPU
 * (src/ggml-cpu/ops.cpp:5121-5165).  dst[i, :] = dequant(table[rows[i], :]).
 * rows_dev: n_rows int32 indices, consecutive indices rows_stride_bytes apart. */
B200_API int b200_get_rows(int type, const void * table_dev, size_t table_row_stride_bytes, int64_t k,
                           int64_t table_rows, const int32_t * rows_dev, size_t rows_stride_bytes, int64_t n_rows,
                           void * dst_dev, int out_type, size_t dst_row_stride_bytes, void * stream);

/* ---- MUL_MAT -------------------------------------------------------------------------------
 * dst[m, n] = W[m, k] (quantized rows) x X[n, k]^T (f32), ggml_mul_mat semantics
 * (include/ggml.h:1287-1293): dst column j (token j) is contiguous over m.
 *
 * b200_mul_mat replaces the dispatcher ggml_cuda_mul_mat (src/ggml-cuda/ggml-cuda.cu:2007-2090):
 *   n <= B200_MMVQ_MAX_N  -> quantize activations to q8_1 + dp4a matvec
 *                            (ggml_cuda_mul_mat_vec_q, src/ggml-cuda/mmvq.cu:500-570)
 *   larger n              -> tensor-core tile kernel (ggml_cuda_mul_mat_q, src/ggml-cuda/mmq.cu:71-143)
 * with ggml's 4-D broadcast rules: ne_b[2] % ne_a[2] == 0, ne_b[3] % ne_a[3] == 0; dims 1..3 of
 * a and b may be permuted/strided (nb arbitrary), dim 0 contiguous; dst is contiguous
 * [ne_a[1], ne_b[1], ne_b[2], ne_b[3]] f32.
 * workspace_dev: at least b200_mul_mat_workspace_size(...) bytes, 256-byte aligned. */
#define B200_MMVQ_MAX_N 8

B200_API size_t b200_mul_mat_workspace_size(int type_a, const int64_t ne_a[4], const int64_t ne_b[4]);
B200_API int    b200_mul_mat(int type_a, const void * a_dev, const int64_t ne_a[4], const size_t nb_a[4],
                             const float * b_dev, const int64_t ne_b[4], const size_t nb_b[4],
                             float * dst_dev, void * workspace_dev, size_t workspace_bytes, void * stream);

/* b200_mul_mat with flags.  B200_MUL_MAT_CONST_WEIGHTS: a_dev is not written by any work queued
 * before this call on `stream` (model weights; ggml marks such buffers
 * GGML_BACKEND_BUFFER_USAGE_WEIGHTS, include/ggml-backend.h:38-42) — the decode kernel may then
 * request weight rows before the preceding kernel has finished (programmatic dependent launch). */
#define B200_MUL_MAT_CONST_WEIGHTS 1u
B200_API int    b200_mul_mat_ex(int type_a, const void * a_dev, const int64_t ne_a[4], const size_t nb_a[4],
                                const float * b_dev, const int64_t ne_b[4], const size_t nb_b[4],
                                float * dst_dev, void * workspace_dev, size_t workspace_bytes,
                                uint32_t flags, void * stream);

/* MUL_MAT_ID (mixture of experts): dst[m, e, t] = as[k, m, ids[e, t]] x b[k, e % ne_b[1], t] for slot e < ne_ids[0], token
 * t < ne_ids[1] = ne_b[2]; as_dev: the experts' weights [k, m, n_as] (quantized rows), ids_dev: int32 [n_used, n_tokens] with
 * byte strides nb_ids (may be a view).  replaces ggml_cuda_mul_mat_id (src/ggml-cuda/ggml-cuda.cu:2092-2233) and the ids
 * path of mul_mat_vec_q (src/ggml-cuda/mmvq.cu:163-166, 556-569); the ids never leave the device.  An id outside
 * [0, n_as) zero-fills its column and raises bit 0 of b200_device_error_flags.  Every (slot, token) pair is a one-column mat-vec
 * on q8_1 activations: Q4_0 / Q8_0 / Q4_K / Q5_K on the TMA-staged kernel (one grid plane per pair, the expert looked up once per
 * CTA), Q6_K and the extra formats on the one-warp-per-row kernel. */
B200_API size_t b200_mul_mat_id_workspace_size(int type_a, const int64_t ne_a[4], const int64_t ne_b[4]);
B200_API int    b200_mul_mat_id(int type_a, const void * as_dev, const int64_t ne_a[4], const size_t nb_a[4],
                                const float * b_dev, const int64_t ne_b[4], const size_t nb_b[4],
                                const int32_t * ids_dev, const int64_t ne_ids[2], const size_t nb_ids[2],
                                float * dst_dev, void * workspace_dev, size_t workspace_bytes, void * stream);

/* A whole list of MUL_MAT nodes in one call, with the semantics of issuing them back to back on `stream` in list order
 * (a node may read what an earlier node of the list writes: decode mat-vecs of one format family then ride ONE
 * persistent launch, chained by per-node completion counters; other hazards cut the list into several launches).
 * replaces the node loop of ggml_backend_cuda_graph_compute -> evaluate_and_capture_cuda_graph ->
 * ggml_cuda_compute_forward (src/ggml-cuda/ggml-cuda.cu:2226-2520, 2870-2960) for graphs made of
 * MUL_MAT nodes: one host call per graph instead of one per node.  workspace_dev is shared by the
 * nodes (stream order makes that safe) and must hold b200_mul_mat_batch_workspace_size() bytes.
 * Returns the first error; nodes before it have been enqueued. */
typedef struct b200_mul_mat_node {
    int32_t       type_a;
    uint32_t      flags;      /* B200_MUL_MAT_* */
    const void  * a_dev;
    int64_t       ne_a[4];
    size_t        nb_a[4];
    const float * b_dev;
    int64_t       ne_b[4];
    size_t        nb_b[4];
    float       * dst_dev;
} b200_mul_mat_node;

B200_API size_t b200_mul_mat_batch_workspace_size(const b200_mul_mat_node * nodes, int64_t count);
B200_API int    b200_mul_mat_batch(const b200_mul_mat_node * nodes, int64_t count,
                                   void * workspace_dev, size_t workspace_bytes, void * stream);

/* ---- row split over the GPUs of one box -------------------------------------------------------
 * replaces the per-MUL_MAT broadcast + peer-copy gather of ggml_cuda_op_mul_mat for split tensors
 * (src/ggml-cuda/ggml-cuda.cu:1441-1759, call sites :1670-1739) and what a separate ncclAllGather would do.
 *
 * Every rank (a process per GPU, or a device of one process) holds its ROW BLOCK of each weight matrix and one copy of
 * a "symmetric region": `bytes` bytes at base[r] for rank r, all copies mapped into this process (CUDA IPC / peer
 * access: NVLink).  The first b200_peer_ctrl_bytes() bytes of the region are control words owned by the library (zero
 * them once, before the first call); the rest is the caller's, laid out identically on every rank.
 * b200_mul_mat_batch_peers is b200_mul_mat_batch with one addition: a node whose dst_dev lies in the LOCAL copy of the
 * region stores every finished output row into EVERY rank's copy at the same offset, from inside the persistent
 * launch, and the call's work on `stream` completes only when the scattered rows of all ranks have landed locally.
 * If rank r passes dst_dev = base[r] + off + 4 * row_low(r) for a node, every rank ends with the whole output vector at
 * base[*] + off -- the all-gather, fused.  A node may read such a gathered vector as its activations (chained inside
 * the launch like any other dependency).  All ranks must issue the same list (SPMD); scattered nodes must be
 * one-column decode nodes of the persistent list kernel (other shapes return B200_ERR_SHAPE).  The caller keeps a rank
 * from starting the next call while another rank still reads the region (as with NCCL). */
typedef struct b200_peer_group {
    int32_t n_ranks, rank;
    void *  base[8];
    size_t  bytes;
} b200_peer_group;
B200_API size_t b200_peer_ctrl_bytes(void);
B200_API int    b200_mul_mat_batch_peers(const b200_mul_mat_node * nodes, int64_t count, const b200_peer_group * peers,
                                         void * workspace_dev, size_t workspace_bytes, void * stream);

/* One-way variant for ONE process driving several devices (ggml's split buffer, src/ggml-cuda/ggml-cuda.cu:1700-1740): every
 * node's finished rows are ALSO stored at dst_dev + remote_delta_bytes, an address in another device's memory (peer
 * access enabled by the caller) -- the row block lands in the main device's dst from inside the launch instead of by a
 * cudaMemcpy2DAsync after it.  No cross-device protocol: the remote rows are complete when this call's work on `stream`
 * is (the caller orders the reader with an event).  One-column decode nodes only. */
B200_API int    b200_mul_mat_batch_remote(const b200_mul_mat_node * nodes, int64_t count, int64_t remote_delta_bytes,
                                          void * workspace_dev, size_t workspace_bytes, void * stream);

/* How the list would be issued, without touching a device: plan[3 i] = launch group of node i (nodes of a group share
 * one launch pair per format family), plan[3 i + 1] = list index of the node it waits for INSIDE its launch (-1: none),
 * plan[3 i + 2] = route (0: per-node call, 1: persistent list kernel, 2: 2..8-column MMA list kernel). */
B200_API int    b200_mul_mat_batch_plan(const b200_mul_mat_node * nodes, int64_t count, int32_t * plan);

/* The two legs separately (2-D, for tests and benches).
 * b200_quantize_acts: f32 [n, k] -> the planar q8 scratch the matvec consumes (same values as
 *   quantize_row_q8_1: int8 plane [n][k] followed by the per-32 scales d as f32 [n][k/32]).
 * b200_mul_mat_vec_q: the matvec proper on that scratch, n <= B200_MMVQ_MAX_N.
 * b200_mul_mat_q: the batched path, any n >= 1 (activations f32, converted inside). */
B200_API size_t b200_acts_scratch_bytes(int64_t k, int64_t n);
B200_API int    b200_quantize_acts(const float * x_dev, size_t x_row_stride_bytes, int64_t k, int64_t n,
                                   void * scratch_dev, void * stream);
B200_API int    b200_mul_mat_vec_q(int type, const void * w_dev, size_t w_row_stride_bytes, int64_t m, int64_t k,
                                   const void * acts_scratch_dev, int64_t n,
                                   float * dst_dev, size_t dst_col_stride_bytes, void * stream);

/* b200_mul_mat_vec_static: b200_quantize_acts + b200_mul_mat_vec_q in one call that allocates nothing — for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It has
 * no counterpart in the reference.  The activation scratch is a
 * static __device__ array of B200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit; a
 * problem with b200_acts_scratch_bytes(k, n) above that returns B200_ERR_TOO_LARGE.  x_dev: f32 [n][k] (rows
 * x_row_stride_bytes apart), n <= B200_MMVQ_MAX_N, dst as for b200_mul_mat_vec_q; results are identical to the
 * two-call sequence.  All calls on a device share the array: calls whose GPU work may overlap (different streams or
 * host threads) race and must be serialized by the caller. */
#define B200_STATIC_SCRATCH_BYTES (1 << 20)
B200_API size_t b200_static_scratch_bytes(void);
B200_API int    b200_mul_mat_vec_static(int type, const void * w_dev, size_t w_row_stride_bytes, int64_t m, int64_t k,
                                        const float * x_dev, size_t x_row_stride_bytes, int64_t n,
                                        float * dst_dev, size_t dst_col_stride_bytes, void * stream);

B200_API size_t b200_mul_mat_q_workspace_size(int64_t m, int64_t k, int64_t n);
B200_API int    b200_mul_mat_q(int type, const void * w_dev, size_t w_row_stride_bytes, int64_t m, int64_t k,
                               const float * x_dev, size_t x_row_stride_bytes, int64_t n,
                               float * dst_dev, size_t dst_col_stride_bytes,
                               void * workspace_dev, size_t workspace_bytes, void * stream);

/* The quantizers of the two list kernels on their own (for bit-exact tests against quantize_row_q8_1_ref,
 * src/ggml-quants.c:225-258).  Layouts (implementation detail of b200_mul_mat_batch's workspace):
 *   record  (decode list kernel, one vector of k values): int8 plane [k], 128-value items with their eight 16-byte
 *            pieces rotated (piece p of item i is stored at piece (p + i) % 8), then f32 d[k/32] (the fp16-rounded scale);
 *   columns (2..8-column MMA list kernel; Q8_K arithmetic, quantize_row_q8_K_ref, src/ggml-quants.c:2555-2592): int8
 *            [k/32][8][32] (the 8 padded columns of a 32-block side by side), then at the next 128-byte boundary f32 d
 *            [k/256][8], then the sums of the 32-blocks as byte pairs (sum & 63, sum >> 6): uint8 [k/256][8][2][2][4] (column,
 *            half of the super-block, low / high bytes, sub-block), then the sums of 16 (bsums) likewise: uint8 [k/256][8][4][2][4]. */
B200_API size_t b200_acts_record_bytes(int64_t k);
B200_API size_t b200_acts_columns_bytes(int64_t k, int64_t n);
B200_API int    b200_quantize_acts_record(const float * x_dev, int64_t k, void * record_dev, void * stream);
B200_API int    b200_quantize_acts_columns(const float * x_dev, size_t x_row_stride_bytes, int64_t k, int64_t n,
                                           void * columns_dev, void * stream);

/* ---- the f32 ops between the MUL_MATs of a decoder layer -----------------------------------------
 * so that a Llama graph stays on the device between our mat-muls.  Values follow the reference's CPU
 * operators (src/ggml-cpu/ops.cpp:4138-4186 rms_norm, :3551-3610 swiglu, :6049-6330 rope, :5685-5790 soft_max;
 * src/ggml-cpu/binary-ops.cpp) within test-backend-ops' NMSE gates; they replace src/ggml-cuda/norm.cu, binbcast.cu,
 * unary.cu (GLU), rope.cu and softmax.cu for these cases.  f32 only; dim 0 contiguous unless stated; dst contiguous.
 *   b200_rms_norm:  dst = x / sqrt(mean(x^2 over dim 0) + eps)
 *   b200_bin_bcast: op 0 = ADD, 1 = MUL; b is broadcast over a (ne_a[i] % ne_b[i] == 0), any strides; dst has a's shape
 *   b200_swiglu:    dst[r, c] = silu(x[r, c]) * g[r, c], rows of nc values (the two halves of one tensor, or two tensors)
 *   b200_rope:      ggml_rope_ext, normal (neox = 0: pairs (2i, 2i+1)) or NeoX (pairs (i, i + n_dims/2)); pos_dev: ne[2] int32;
 *                   freq_factors_dev: n_dims/2 floats or NULL; corr_dim0/1 from ggml_rope_yarn_corr_dims (src/ggml.c:4088-4096)
 *   b200_soft_max:  softmax over dim 0 of x * scale + slope(head) * mask, mask f32 / f16 [ne0, >= ne1, ne2m, ne3m] broadcast over
 *                   dims 2, 3 (or NULL), ALiBi slopes from max_bias (head = dim 2), optional per-head sinks */
B200_API int b200_rms_norm(const float * x_dev, const int64_t ne[4], const size_t nb[4], float * dst_dev, float eps, void * stream);
B200_API int b200_bin_bcast(int op, const float * a_dev, const int64_t ne_a[4], const size_t nb_a[4],
                            const float * b_dev, const int64_t ne_b[4], const size_t nb_b[4], float * dst_dev, void * stream);
B200_API int b200_swiglu(const float * x_dev, size_t x_row_stride_bytes, const float * g_dev, size_t g_row_stride_bytes,
                         int64_t nc, int64_t rows, float * dst_dev, void * stream);
B200_API int b200_rope(const float * x_dev, const int64_t ne[4], const size_t nb[4], float * dst_dev, const size_t nb_dst[4],
                       const int32_t * pos_dev, const float * freq_factors_dev, int n_dims, int neox,
                       float freq_base, float freq_scale, float ext_factor, float attn_factor, float corr_dim0, float corr_dim1, void * stream);
B200_API int b200_soft_max(const float * x_dev, const int64_t ne[4], const size_t nb[4], const void * mask_dev, int mask_type,
                           const int64_t ne_mask[4], const size_t nb_mask[4], const float * sinks_dev, float scale, float max_bias,
                           float * dst_dev, void * stream);

/* ---- GGUF file -> device ---------------------------------------------------------------------
 * The container on the input side of the path.  b200_gguf_open reads header, key/value pairs and tensor
 * infos with the acceptance rules of gguf_init_from_file_impl (src/gguf.cpp:319-735; format include/gguf.h:1-31) and
 * keeps the file open; no tensor data is read.  replaces gguf_init_from_file(no_alloc) + the gguf_get_* getters
 * (include/gguf.h:78-128) for reading, and — b200_gguf_load — the per-tensor read + ggml_backend_tensor_set loop of the
 * reference's examples (a synchronous pageable cudaMemcpy per tensor, src/ggml-cuda/ggml-cuda.cu:614-621): tensor
 * bytes go file -> a ring of eight 16 MiB pinned staging buffers -> device with cudaMemcpyAsync, one reader
 * thread per stage filling the next stages while earlier ones are in flight.  The bytes on the device are ggml's canonical block bytes.
 * Host-only calls (everything but the two load calls) need no GPU.  Value types = enum gguf_type (include/gguf.h:53-68):
 * 0 u8, 1 i8, 2 u16, 3 i16, 4 u32, 5 i32, 6 f32, 7 bool, 8 string, 9 array, 10 u64, 11 i64, 12 f64. */
typedef struct b200_gguf b200_gguf;
typedef struct b200_gguf_tensor_info {
    const char * name;          /* owned by the handle */
    int32_t      type;          /* enum ggml_type */
    int32_t      n_dims;
    int64_t      ne[4];
    uint64_t     file_offset;   /* absolute byte offset of the tensor's data in the file */
    uint64_t     nbytes;        /* ggml_nbytes */
    uint64_t     device_offset; /* where b200_gguf_load puts it, relative to base_dev (256-byte aligned) */
} b200_gguf_tensor_info;
B200_API int          b200_gguf_open(const char * path, b200_gguf ** out);      /* B200_ERR_FILE + b200_gguf_last_error() */
B200_API void         b200_gguf_close(b200_gguf * g);
B200_API const char * b200_gguf_last_error(void);                               /* of the calling thread */
B200_API uint32_t     b200_gguf_version(const b200_gguf * g);
B200_API size_t       b200_gguf_alignment(const b200_gguf * g);
B200_API uint64_t     b200_gguf_data_offset(const b200_gguf * g);
B200_API uint64_t     b200_gguf_data_bytes(const b200_gguf * g);
B200_API int64_t      b200_gguf_n_kv(const b200_gguf * g);
B200_API int64_t      b200_gguf_find_key(const b200_gguf * g, const char * key);          /* -1 if absent */
B200_API const char * b200_gguf_key(const b200_gguf * g, int64_t key_id);
B200_API int          b200_gguf_kv_type(const b200_gguf * g, int64_t key_id);             /* 9 for arrays */
B200_API int          b200_gguf_arr_type(const b200_gguf * g, int64_t key_id);
B200_API uint64_t     b200_gguf_arr_n(const b200_gguf * g, int64_t key_id);
B200_API const void * b200_gguf_val_data(const b200_gguf * g, int64_t key_id);            /* raw little-endian scalar / array bytes */
B200_API const char * b200_gguf_val_str(const b200_gguf * g, int64_t key_id, uint64_t j); /* string value (j = 0) or j-th array string */
B200_API int          b200_gguf_val_number(const b200_gguf * g, int64_t key_id, double * as_f64, int64_t * as_i64);
B200_API int64_t      b200_gguf_n_tensors(const b200_gguf * g);
B200_API int64_t      b200_gguf_find_tensor(const b200_gguf * g, const char * name);      /* -1 if absent */
B200_API int          b200_gguf_tensor(const b200_gguf * g, int64_t tensor_id, b200_gguf_tensor_info * out);
B200_API size_t       b200_gguf_type_row_bytes(int type, int64_t ne0);                    /* ggml_row_size for every ggml type; 0 = invalid */
/* Device image: every tensor at base_dev + device_offset, in file order.  base_dev 256-byte aligned, bytes >=
 * b200_gguf_device_bytes().  The call returns when the last copy has left the staging buffers (the handle may then be
 * closed); the copies are ordered on `stream` like any cudaMemcpyAsync. */
B200_API size_t       b200_gguf_device_bytes(const b200_gguf * g);
B200_API int          b200_gguf_load(b200_gguf * g, void * base_dev, size_t bytes, void * stream);
/* The same for chosen tensors at caller-chosen addresses (the plugin: tensor->data of a ggml buffer). */
B200_API int          b200_gguf_load_tensors(b200_gguf * g, const int64_t * tensor_ids, void * const * dst_dev, int64_t count, void * stream);

/* ---- what the tensor cores of this box deliver -------------------------------------------------
 * A tcgen05 micro-kernel (operands resident in shared memory, accumulator in TMEM, nothing but the MMA issue loop):
 * kind 0 = kind::i8 (s8 x s8 -> s32), kind 1 = kind::f16 with bf16 operands (-> f32).  The denominators of the batched
 * MUL_MAT's roofline (BASELINE.json: "int8 TOPS ... against the int8 dense peak"); the reference measures no peak.
 * a_dev: [128][128] bytes, b_dev: [256][128] bytes, row-major, K contiguous.  out_dev != NULL: one CTA, one pass over
 * the 128 bytes of K, accumulators written as [128][256] 32-bit words (verification).  out_dev == NULL: every SM issues
 * iters x 4 MMAs of 128 x 256 x 32 bytes per launch, `reps` launches are timed with CUDA events (the call synchronizes
 * on its own events) and 2 x MACs / s over the whole GPU is returned in *ops_per_s. */
B200_API int b200_tensor_peak(int kind, const void * a_dev, const void * b_dev, void * out_dev, int iters, int reps,
                              double * ops_per_s, void * stream);

/* Number of kernels this library has launched since load (bench.py's gpu_launches). */
B200_API uint64_t b200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_H */
