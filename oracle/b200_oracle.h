/*
 * b200_oracle.h — CPU ORACLE, TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the reference's CPU algorithms for the hot path (ggml block formats
 * Q4_0 / Q8_0 / Q4_K / Q5_K / Q6_K: dequantization, q8_1 activation quantization, GET_ROWS and the
 * MUL_MAT contraction).  It is the checker the CUDA path is compared against; it is never the thing
 * measured or shipped.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  The product library (libggml_b200_kernels.so /
 * libggml-b200.so) does not link, import or call anything in this directory.
 *
 * PARITY PINNING: every function here is checked bit-for-bit against the unmodified reference
 * compiled from /root/reference into oracle/_ref/ (tests/test_oracle_pinning.py) and against the
 * committed reference-generated golden vectors in tests/golden/ (oracle/gen_golden.py).
 *
 * All citations are relative to /root/reference.
 */
#ifndef B200_ORACLE_H
#define B200_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* type ids: the numeric values of the reference's enum ggml_type (include/ggml.h:376-410) */
enum b200o_type {
    B200O_F32  = 0,
    B200O_F16  = 1,
    B200O_Q4_0 = 2,
    B200O_Q4_1 = 3,
    B200O_Q5_0 = 6,
    B200O_Q5_1 = 7,
    B200O_Q8_0 = 8,
    B200O_Q8_1 = 9,
    B200O_Q4_K = 12,
    B200O_Q5_K = 13,
    B200O_Q6_K = 14,
    B200O_IQ4_NL = 20,
};

/* elements per block / bytes per block (src/ggml-common.h:170-336) */
int    b200o_block_elems(int type);
size_t b200o_block_bytes(int type);
size_t b200o_row_bytes(int type, int64_t k);

/* IEEE binary16 <-> binary32, round-to-nearest-even, NaN -> sign|0x7E00
 * (same values as ggml_compute_fp32_to_fp16 / ggml_compute_fp16_to_fp32, src/ggml-impl.h:358-405) */
float    b200o_fp16_to_fp32(uint16_t h);
uint16_t b200o_fp32_to_fp16(float f);

/* dequantize_row_<type>: k elements (k % block_elems == 0) from blocks at x -> f32 at y.
 * src/ggml-quants.c:307-325 (q4_0), 401-415 (q8_0), 1352-1374 (q4_K), 1554-1579 (q5_K), 1762-1791 (q6_K),
 * 327-346 (q4_1), 348-372 (q5_0), 374-399 (q5_1), 2512-2528 (iq4_nl) */
void b200o_dequantize_row(int type, const void * x, float * y, int64_t k);
/* same, then f32 -> f16 RN-even (what "dequantize to f16" means for the CPU path: to_float then
 * ggml_fp32_to_fp16_row, src/ggml.c:455-470) */
void b200o_dequantize_row_f16(int type, const void * x, uint16_t * y, int64_t k);

/* quantize_row_q8_1_ref (src/ggml-quants.c:225-258): f32[k] -> block_q8_1[k/32] */
void b200o_quantize_row_q8_1(const float * x, void * y, int64_t k);
/* quantize_row_q8_K_ref (src/ggml-quants.c:2555-2592): f32[k], k % 256 == 0 -> planar q[k] int8, d[k/256] f32, bsums[k/16]
 * int16: the activation format of ggml-cpu's K-quant MUL_MAT (and of this backend's 2..8-column decode kernel) */
void b200o_quantize_row_q8_K(const float * x, int8_t * q, float * d, int16_t * bsums, int64_t k);
/* quantize_row_q8_0_ref (src/ggml-quants.c:199-222) and quantize_row_q4_0_ref (:41-76): used to
 * make random-quantized weights without the reference present */
void b200o_quantize_row_q8_0(const float * x, void * y, int64_t k);
void b200o_quantize_row_q4_0(const float * x, void * y, int64_t k);

/* GET_ROWS on a quantized table (src/ggml-cpu/ops.cpp:5121-5165 -> to_float per selected row):
 * dst[i, :] = dequant(table[rows[i], :]) ; table rows are row_stride bytes apart */
void b200o_get_rows(int type, const void * table, size_t row_stride, int64_t k,
                    const int32_t * rows, int64_t nrows_out, float * dst);

/* MUL_MAT, dst[j*m + i] = sum_k W[i,k] * X[j,k]  (include/ggml.h:1287-1293; dst is [m, n], column j
 * = token j, contiguous over m).  Two definitions of truth, both accumulated in double:
 *   mode 0: W dequantized, X exact f32          (what an infinitely precise backend would give)
 *   mode 1: W dequantized, X round-tripped through q8_1 (quantize_row_q8_1_ref then d*q) — the
 *           arithmetic the GPU path (and the reference CUDA path, src/ggml-cuda/mmvq.cu:536-543)
 *           performs; the GPU result must match this one to f32 rounding.
 *   mode 2: W dequantized, X round-tripped through Q8_K (quantize_row_q8_K_ref then d*q; k % 256 == 0) — the arithmetic of
 *           ggml-cpu's K-quant MUL_MAT (src/ggml-cpu/quants.c:550-623 et al.) and of this backend's decode kernel for
 *           2..8 columns of Q4_K / Q5_K / Q6_K weights.
 * W rows are row_stride bytes apart; X rows are k floats. */
void b200o_mul_mat(int type, const void * w, size_t row_stride, int64_t m, int64_t k,
                   const float * x, int64_t n, float * dst, int mode);

/* normalized mean squared error as test-backend-ops computes it (tests/test-backend-ops.cpp:283-296) */
double b200o_nmse(const float * a, const float * b, size_t n);

#ifdef __cplusplus
}
#endif
#endif /* B200_ORACLE_H */
