// dequant.cu — dequantize_row_* (f32 / f16 out), the strided CPY q->f form, and GET_ROWS on quantized
// tables, for Q4_0 / Q8_0 / Q4_K / Q5_K / Q6_K.
//
// Value contract: bit-exact with the reference CPU dequantizers (src/ggml-quants.c:307-325,
// 401-415, 1352-1374, 1554-1579, 1762-1791).  Those are compiled without FMA contraction, so every
// product and difference here is an explicit round-to-nearest intrinsic (__fmul_rn / __fsub_rn):
// the compiler can never fuse them, whatever flags the TU is built with.
//
// Shape of the work: HBM-bound on the OUTPUT side (0.56-1.06 B in, 4 B out per element).  One thread
// produces 8 consecutive elements of one block (one 32-byte or 16-byte vector store); a warp writes
// 1 KiB of contiguous f32.  Grid = ceil(n_elems / 8 / 256) CTAs of 256 threads.
#include "common.cuh"

namespace b200 {

// ---------------------------------------------------------------- 8 consecutive elements of a block
// c = index of the 8-element chunk inside the block (0..3 for 32-blocks, 0..31 for 256-blocks).
template <int T> __device__ __forceinline__ void dequant8(const uint8_t * b, int c, float (&y)[8]);

template <> __device__ __forceinline__ void dequant8<T_Q4_0>(const uint8_t * b, int c, float (&y)[8]) {
    const float d = f16_bits_to_f32(ld_u16(b));
    // elements 0..15 = low nibbles of qs[0..15], 16..31 = high nibbles of qs[0..15]
    const uint8_t * q = b + 2 + 8 * (c & 1);
    const uint32_t w0 = ld_u32_a2(q), w1 = ld_u32_a2(q + 4);
    const int sh = (c >> 1) * 4;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t w = i < 4 ? w0 : w1;
        const int v = (int) ((w >> (8 * (i & 3) + sh)) & 0xFu) - 8;
        y[i] = __fmul_rn((float) v, d);
    }
}

template <> __device__ __forceinline__ void dequant8<T_Q8_0>(const uint8_t * b, int c, float (&y)[8]) {
    const float d = f16_bits_to_f32(ld_u16(b));
    const uint8_t * q = b + 2 + 8 * c;
    const uint32_t w0 = ld_u32_a2(q), w1 = ld_u32_a2(q + 4);
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t w = i < 4 ? w0 : w1;
        const int v = (int) (int8_t) ((w >> (8 * (i & 3))) & 0xFFu);
        y[i] = __fmul_rn((float) v, d);
    }
}

// ---- the formats only the general kernels serve (32-weight blocks, 2-byte aligned; reference src/ggml-quants.c:327-399, 2512-2528).
// Chunk c: elements 8c..8c+7; elements 0..15 are the low nibbles of qs[0..15], 16..31 the high nibbles.
template <bool HAS_M, bool HAS_QH, bool NL>
__device__ __forceinline__ void dequant8_32(const uint8_t * b, int c, float (&y)[8]) {
    const float d = f16_bits_to_f32(ld_u16(b));
    const float m = HAS_M ? f16_bits_to_f32(ld_u16(b + 2)) : 0.0f;
    constexpr int oqh = HAS_M ? 4 : 2, oqs = oqh + (HAS_QH ? 4 : 0);
    const uint8_t * q = b + oqs + 8 * (c & 1);
    const uint32_t w0 = ld_u32_a2(q), w1 = ld_u32_a2(q + 4);
    uint32_t qh = 0;
    if (HAS_QH) qh = ld_u32_a2(b + oqh) >> (8 * c);          // bit j (low nibbles) / j + 16 (high nibbles): element 8c + i -> bit 8c + i
    const int sh = (c >> 1) * 4;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t w = i < 4 ? w0 : w1;
        int v = (int) ((w >> (8 * (i & 3) + sh)) & 0xFu);
        if (HAS_QH) v |= (int) ((qh >> i) & 1u) << 4;
        if (NL)               y[i] = __fmul_rn(d, (float) iq4nl_value(v));                       // d * kvalues_iq4nl[q]
        else if (HAS_M)       y[i] = __fadd_rn(__fmul_rn((float) v, d), m);                      // x * d + m
        else if (HAS_QH)      y[i] = __fmul_rn((float) (v - 16), d);                             // (x - 16) * d
    }
}
template <> __device__ __forceinline__ void dequant8<T_Q4_1>(const uint8_t * b, int c, float (&y)[8]) { dequant8_32<true, false, false>(b, c, y); }
template <> __device__ __forceinline__ void dequant8<T_Q5_0>(const uint8_t * b, int c, float (&y)[8]) { dequant8_32<false, true, false>(b, c, y); }
template <> __device__ __forceinline__ void dequant8<T_Q5_1>(const uint8_t * b, int c, float (&y)[8]) { dequant8_32<true, true, false>(b, c, y); }
template <> __device__ __forceinline__ void dequant8<T_IQ4_NL>(const uint8_t * b, int c, float (&y)[8]) { dequant8_32<false, false, true>(b, c, y); }

// Q4_K / Q5_K share everything but the 5th bit.  Element e = 8c: 64-element group g = e/64; inside
// the group the first 32 are low nibbles (scale pair 2g), the next 32 high nibbles (pair 2g+1) of
// qs[32g .. 32g+31].
template <bool HAS_QH> __device__ __forceinline__ void dequant8_k45(const uint8_t * b, int c, float (&y)[8]) {
    // blocks are 144 / 176 bytes: every block of a 16-byte aligned tensor is 16-byte aligned -> one 16-byte load for
    // the header, one 8-byte load per quant run (the 2-byte-load path costs 12-16 loads per 8 elements and is
    // load-issue bound: 3.6 TB/s, profiles/r01i_support_probe.txt)
    const bool al = (reinterpret_cast<uintptr_t>(b) & 15) == 0;
    uint32_t dm, s0, s1, s2;
    if (al) { const uint4 hd = ld_u128(b); dm = hd.x; s0 = hd.y; s1 = hd.z; s2 = hd.w; }
    else    { dm = ld_u32_a2(b); s0 = ld_u32_a2(b + 4); s1 = ld_u32_a2(b + 8); s2 = ld_u32_a2(b + 12); }
    const float d    = f16_bits_to_f32(dm);
    const float dmin = f16_bits_to_f32(dm >> 16);
    const int g  = c >> 3;            // 64-element group
    const int hi = (c >> 2) & 1;      // 0: low nibbles, 1: high nibbles
    const int l0 = (c & 3) * 8;       // first byte inside the group's 32 bytes
    int sc, mn;
    k4_scale_min(s0, s1, s2, 2 * g + hi, sc, mn);
    const float d1 = __fmul_rn(d, (float) sc);
    const float m1 = __fmul_rn(dmin, (float) mn);
    const uint8_t * q = b + (HAS_QH ? 48 : 16) + 32 * g + l0;
    uint32_t w0, w1, h0 = 0, h1 = 0;
    if (al) {
        const uint2 w = ld_u64(q); w0 = w.x; w1 = w.y;
        if (HAS_QH) { const uint2 h = ld_u64(b + 16 + l0); h0 = h.x; h1 = h.y; }
    } else {
        w0 = ld_u32_a2(q); w1 = ld_u32_a2(q + 4);            // (never past the block: the last run ends at its last byte)
        if (HAS_QH) { h0 = ld_u32_a2(b + 16 + l0); h1 = ld_u32_a2(b + 16 + l0 + 4); }
    }
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t w = i < 4 ? w0 : w1;
        int v = (int) ((w >> (8 * (i & 3) + 4 * hi)) & 0xFu);
        if (HAS_QH) {
            const uint32_t h = i < 4 ? h0 : h1;
            v += (int) ((h >> (8 * (i & 3) + 2 * g + hi)) & 1u) << 4;
        }
        y[i] = __fsub_rn(__fmul_rn(d1, (float) v), m1);
    }
}
template <> __device__ __forceinline__ void dequant8<T_Q4_K>(const uint8_t * b, int c, float (&y)[8]) { dequant8_k45<false>(b, c, y); }
template <> __device__ __forceinline__ void dequant8<T_Q5_K>(const uint8_t * b, int c, float (&y)[8]) { dequant8_k45<true>(b, c, y); }

// Q6_K: element e = 128h + 32s + l (h half, s stream 0..3, l 0..31):
//   low 4 bits = ql[64h + 32(s&1) + l] nibble (s>>1); high 2 bits = (qh[32h + l] >> 2s) & 3;
//   scale = scales[8h + 2s + l/16]; value = (d*scale) * (q - 32), two rounded products.
template <> __device__ __forceinline__ void dequant8<T_Q6_K>(const uint8_t * b, int c, float (&y)[8]) {
    const int h  = c >> 4;
    const int s  = (c >> 2) & 3;
    const int l0 = (c & 3) * 8;
    const float d = f16_bits_to_f32(ld_u16(b + 208));
    const int   sc = (int) (int8_t) __ldg(b + 192 + 8 * h + 2 * s + (l0 >> 4));
    const float a  = __fmul_rn(d, (float) sc);
    const uint8_t * ql = b + 64 * h + 32 * (s & 1) + l0;
    const uint8_t * qh = b + 128 + 32 * h + l0;
    uint32_t w0, w1, h0, h1;
    ld_2x32_a2(ql, w0, w1);           // a warp covers one block: the alignment branch is warp-uniform
    ld_2x32_a2(qh, h0, h1);
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t w = i < 4 ? w0 : w1;
        const uint32_t hh = i < 4 ? h0 : h1;
        const int lo = (int) ((w  >> (8 * (i & 3) + 4 * (s >> 1))) & 0xFu);
        const int hi = (int) ((hh >> (8 * (i & 3) + 2 * s)) & 3u);
        const int q  = (lo | (hi << 4)) - 32;
        y[i] = __fmul_rn(a, (float) q);
    }
}

// ---------------------------------------------------------------- stores
template <typename OUT> __device__ __forceinline__ void store8(OUT * p, const float (&y)[8], bool vec_ok);

template <> __device__ __forceinline__ void store8<float>(float * p, const float (&y)[8], bool vec_ok) {
    if (vec_ok) {
        reinterpret_cast<float4 *>(p)[0] = make_float4(y[0], y[1], y[2], y[3]);
        reinterpret_cast<float4 *>(p)[1] = make_float4(y[4], y[5], y[6], y[7]);
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) p[i] = y[i];
    }
}
template <> __device__ __forceinline__ void store8<__half>(__half * p, const float (&y)[8], bool vec_ok) {
    if (vec_ok) {
        uint4 v;
        __half2 t;
        t = __halves2half2(__float2half_rn(y[0]), __float2half_rn(y[1])); v.x = *reinterpret_cast<uint32_t *>(&t);
        t = __halves2half2(__float2half_rn(y[2]), __float2half_rn(y[3])); v.y = *reinterpret_cast<uint32_t *>(&t);
        t = __halves2half2(__float2half_rn(y[4]), __float2half_rn(y[5])); v.z = *reinterpret_cast<uint32_t *>(&t);
        t = __halves2half2(__float2half_rn(y[6]), __float2half_rn(y[7])); v.w = *reinterpret_cast<uint32_t *>(&t);
        *reinterpret_cast<uint4 *>(p) = v;
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) p[i] = __float2half_rn(y[i]);
    }
}

// f32 output, a full warp of consecutive 8-element chunks: a lane's 32 bytes would go out as two 16-byte stores that each
// fill HALF of a 32-byte sector.  Lane pairs swap halves first, so that every store instruction of the warp writes whole
// sectors (store 1: even lane its own first half at +0, odd lane the even lane's second half at +16; store 2: even lane the
// odd lane's first half at +32, odd lane its own second half at +48 -- branch-free: both lanes of a pair are in the same
// instruction).  p: this lane's chunk, 32-byte aligned.  Measured on a 32000 x 4096 table: Q4_0 5.1 -> 6.0, Q8_0 5.7 -> 6.5,
// Q4_K 4.7 -> 5.2 TB/s (profiles/r03n_support_probe.txt).
__device__ __forceinline__ void store8_paired(float * p, const float (&v)[8]) {
    const bool odd = threadIdx.x & 1;
    float t[4];
#pragma unroll
    for (int j = 0; j < 4; j++) t[j] = __shfl_xor_sync(0xFFFFFFFFu, odd ? v[j] : v[4 + j], 1);
    float * q = p - (odd ? 8 : 0);                                      // the pair's 64 bytes start at the even lane's chunk
    const float4 s1 = odd ? make_float4(t[0], t[1], t[2], t[3]) : make_float4(v[0], v[1], v[2], v[3]);
    const float4 s2 = odd ? make_float4(v[4], v[5], v[6], v[7]) : make_float4(t[0], t[1], t[2], t[3]);
    reinterpret_cast<float4 *>(q)[odd ? 1 : 0] = s1;
    reinterpret_cast<float4 *>(q)[odd ? 3 : 2] = s2;
}

// bf16 output (replaces ggml_get_to_bf16_cuda, src/ggml-cuda/convert.cu:650-659): RN-even narrowing of the f32 values
template <> __device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16 * p, const float (&y)[8], bool vec_ok) {
    if (vec_ok) {
        uint4 v;
        __nv_bfloat162 t;
        t = __floats2bfloat162_rn(y[0], y[1]); v.x = *reinterpret_cast<uint32_t *>(&t);
        t = __floats2bfloat162_rn(y[2], y[3]); v.y = *reinterpret_cast<uint32_t *>(&t);
        t = __floats2bfloat162_rn(y[4], y[5]); v.z = *reinterpret_cast<uint32_t *>(&t);
        t = __floats2bfloat162_rn(y[6], y[7]); v.w = *reinterpret_cast<uint32_t *>(&t);
        *reinterpret_cast<uint4 *>(p) = v;
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) p[i] = __float2bfloat16_rn(y[i]);
    }
}

// ---------------------------------------------------------------- contiguous row(s)
template <int T, typename OUT>
__global__ void __launch_bounds__(256) k_dequantize_row(const uint8_t * __restrict__ x, OUT * __restrict__ y, int64_t n_chunks) {
    const int64_t i = (int64_t) blockIdx.x * 256 + threadIdx.x;   // 8-element chunk index
    const bool full_warp = __all_sync(0xFFFFFFFFu, i < n_chunks);
    if (i >= n_chunks) return;
    constexpr int cpb = fmt<T>::qk / 8;
    const int64_t blk = i / cpb;
    const int     c   = (int) (i - blk * cpb);
    float v[8];
    dequant8<T>(x + blk * fmt<T>::bytes, c, v);
    OUT * p = y + i * 8;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(y) & (8 * sizeof(OUT) - 1)) == 0;      // warp-uniform
    if (sizeof(OUT) == 4 && full_warp && vec_ok) { store8_paired(reinterpret_cast<float *>(p), v); return; }
    store8<OUT>(p, v, vec_ok);
}

// ---------------------------------------------------------------- 4-D strided (CPY / DUP / CONT q -> f)
struct nd4 { int64_t ne[4]; int64_t nb[4]; };

template <int T, typename OUT>
__global__ void __launch_bounds__(256) k_dequantize_4d(const uint8_t * __restrict__ x, uint8_t * __restrict__ y,
                                                       nd4 src, nd4 dst, int64_t n_chunks, int same_rows) {
    const int64_t i = (int64_t) blockIdx.x * 256 + threadIdx.x;
    if (i >= n_chunks) return;
    // source coordinates of logical element 8*i
    const int64_t cpr = src.ne[0] / 8;                       // chunks per source row
    int64_t r = i / cpr;
    const int64_t e0 = (i - r * cpr) * 8;
    const int64_t s1 = r % src.ne[1]; r /= src.ne[1];
    const int64_t s2 = r % src.ne[2];
    const int64_t s3 = r / src.ne[2];
    const uint8_t * row = x + s1 * src.nb[1] + s2 * src.nb[2] + s3 * src.nb[3];
    const int64_t blk = e0 / fmt<T>::qk;
    float v[8];
    dequant8<T>(row + blk * fmt<T>::bytes, (int) ((e0 - blk * fmt<T>::qk) / 8), v);

    if (same_rows) {
        // 8 consecutive logical elements stay inside one destination row
        int64_t lin = i * 8;
        const int64_t d0 = lin % dst.ne[0]; lin /= dst.ne[0];
        const int64_t d1 = lin % dst.ne[1]; lin /= dst.ne[1];
        const int64_t d2 = lin % dst.ne[2];
        const int64_t d3 = lin / dst.ne[2];
        uint8_t * p = y + d0 * dst.nb[0] + d1 * dst.nb[1] + d2 * dst.nb[2] + d3 * dst.nb[3];
        if (dst.nb[0] == (int64_t) sizeof(OUT)) {
            store8<OUT>(reinterpret_cast<OUT *>(p), v, (reinterpret_cast<uintptr_t>(p) & (8 * sizeof(OUT) - 1)) == 0);
        } else {
#pragma unroll
            for (int j = 0; j < 8; j++) {
                if (sizeof(OUT) == 4) *reinterpret_cast<float *>(p + j * dst.nb[0]) = v[j];
                else                  *reinterpret_cast<__half *>(p + j * dst.nb[0]) = __float2half_rn(v[j]);
            }
        }
    } else {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            int64_t lin = i * 8 + j;
            const int64_t d0 = lin % dst.ne[0]; lin /= dst.ne[0];
            const int64_t d1 = lin % dst.ne[1]; lin /= dst.ne[1];
            const int64_t d2 = lin % dst.ne[2];
            const int64_t d3 = lin / dst.ne[2];
            uint8_t * p = y + d0 * dst.nb[0] + d1 * dst.nb[1] + d2 * dst.nb[2] + d3 * dst.nb[3];
            if (sizeof(OUT) == 4) *reinterpret_cast<float *>(p) = v[j];
            else                  *reinterpret_cast<__half *>(p) = __float2half_rn(v[j]);
        }
    }
}

// ---------------------------------------------------------------- GET_ROWS
// grid (ceil(k/8/256), n_rows); a row index outside the table aborts in the reference (GGML_ASSERT,
// src/ggml-cpu/ops.cpp:5147).  A kernel cannot return a code: such rows are ZERO-FILLED (never stale memory)
// and bit 0 of the library's sticky error word is set (b200_device_error_flags -> B200_ERR_INDEX; the plugin
// aborts on it at its next synchronize, like the reference).
template <int T, typename OUT>
__global__ void __launch_bounds__(256) k_get_rows(const uint8_t * __restrict__ table, int64_t row_stride, int64_t k, int64_t table_rows,
                                                  const uint8_t * __restrict__ rows, int64_t rows_stride,
                                                  uint8_t * __restrict__ dst, int64_t dst_row_stride, unsigned * __restrict__ bad) {
    const int64_t i = (int64_t) blockIdx.x * 256 + threadIdx.x;
    const bool full_warp = __all_sync(0xFFFFFFFFu, i * 8 < k);
    if (i * 8 >= k) return;
    const int64_t out_row = blockIdx.y;
    const int32_t r = *reinterpret_cast<const int32_t *>(rows + out_row * rows_stride);
    OUT * p = reinterpret_cast<OUT *>(dst + out_row * dst_row_stride) + i * 8;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(p) & (8 * sizeof(OUT) - 1)) == 0;
    // (r and the row base are warp-uniform, and with them every branch below)
    const bool paired = sizeof(OUT) == 4 && full_warp && ((reinterpret_cast<uintptr_t>(dst + out_row * dst_row_stride) & 31) == 0);
    float v[8];
    if (r < 0 || r >= table_rows) {
        if (bad && i == 0) atomicOr_system(bad, 1u);
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = 0.0f;
    } else {
        constexpr int cpb = fmt<T>::qk / 8;
        const int64_t blk = i / cpb;
        dequant8<T>(table + (int64_t) r * row_stride + blk * fmt<T>::bytes, (int) (i - blk * cpb), v);
    }
    if (paired) store8_paired(reinterpret_cast<float *>(p), v);
    else        store8<OUT>(p, v, vec_ok);
}

} // namespace b200

// ================================================================= C ABI
using namespace b200;


#define DISPATCH_OUT(T, out_type, CALL) \
    if (out_type == T_F32) { CALL(T, float) } else if (out_type == T_F16) { CALL(T, __half) } else { CALL(T, __nv_bfloat16) }
#define DISPATCH_TYPE_OUT(type, out_type, CALL)                                              \
    switch (type) {                                                                          \
        case T_Q4_0: DISPATCH_OUT(T_Q4_0, out_type, CALL) break;                             \
        case T_Q8_0: DISPATCH_OUT(T_Q8_0, out_type, CALL) break;                             \
        case T_Q4_K: DISPATCH_OUT(T_Q4_K, out_type, CALL) break;                             \
        case T_Q5_K: DISPATCH_OUT(T_Q5_K, out_type, CALL) break;                             \
        case T_Q6_K: DISPATCH_OUT(T_Q6_K, out_type, CALL) break;                             \
        case T_Q4_1: DISPATCH_OUT(T_Q4_1, out_type, CALL) break;                             \
        case T_Q5_0: DISPATCH_OUT(T_Q5_0, out_type, CALL) break;                             \
        case T_Q5_1: DISPATCH_OUT(T_Q5_1, out_type, CALL) break;                             \
        case T_IQ4_NL: DISPATCH_OUT(T_IQ4_NL, out_type, CALL) break;                         \
        default: return B200_ERR_TYPE;                                                       \
    }

extern "C" int b200_dequantize_row(int type, const void * x_dev, void * y_dev, int64_t n_elems, int out_type, void * stream) {
    if (!type_is_block(type) || (out_type != T_F32 && out_type != T_F16 && out_type != T_BF16)) return B200_ERR_TYPE;
    if (n_elems < 0 || n_elems % type_qk(type) != 0) return B200_ERR_SHAPE;
    if (n_elems == 0) return B200_OK;
    const int64_t n_chunks = n_elems / 8;
    const dim3 grid((unsigned) ceil_div(n_chunks, 256));
    cudaStream_t st = (cudaStream_t) stream;
#define CALL(T, OUT) k_dequantize_row<T, OUT><<<grid, 256, 0, st>>>((const uint8_t *) x_dev, (OUT *) y_dev, n_chunks);
    DISPATCH_TYPE_OUT(type, out_type, CALL)
#undef CALL
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

extern "C" int b200_dequantize_4d(int type, const void * x_dev, const int64_t ne_src[4], const size_t nb_src[4],
                                  void * y_dev, int out_type, const int64_t ne_dst[4], const size_t nb_dst[4], void * stream) {
    if (!type_is_block(type) || (out_type != T_F32 && out_type != T_F16 && out_type != T_BF16)) return B200_ERR_TYPE;
    int64_t n_src = 1, n_dst = 1;
    for (int i = 0; i < 4; i++) {
        if (ne_src[i] < 0 || ne_dst[i] < 0) return B200_ERR_SHAPE;
        n_src *= ne_src[i]; n_dst *= ne_dst[i];
    }
    if (n_src != n_dst || ne_src[0] % type_qk(type) != 0) return B200_ERR_SHAPE;
    if (nb_src[0] != (size_t) type_bytes(type)) return B200_ERR_STRIDE;
    if (n_src == 0) return B200_OK;
    nd4 s, d;
    for (int i = 0; i < 4; i++) { s.ne[i] = ne_src[i]; s.nb[i] = (int64_t) nb_src[i]; d.ne[i] = ne_dst[i]; d.nb[i] = (int64_t) nb_dst[i]; }
    const int same_rows = (ne_dst[0] % 8 == 0) ? 1 : 0;
    const int64_t n_chunks = n_src / 8;
    const dim3 grid((unsigned) ceil_div(n_chunks, 256));
    cudaStream_t st = (cudaStream_t) stream;
#define CALL(T, OUT) k_dequantize_4d<T, OUT><<<grid, 256, 0, st>>>((const uint8_t *) x_dev, (uint8_t *) y_dev, s, d, n_chunks, same_rows);
    DISPATCH_TYPE_OUT(type, out_type, CALL)
#undef CALL
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

extern "C" int b200_get_rows(int type, const void * table_dev, size_t table_row_stride_bytes, int64_t k, int64_t table_rows,
                             const int32_t * rows_dev, size_t rows_stride_bytes, int64_t n_rows,
                             void * dst_dev, int out_type, size_t dst_row_stride_bytes, void * stream) {
    if (!type_is_block(type) || (out_type != T_F32 && out_type != T_F16 && out_type != T_BF16)) return B200_ERR_TYPE;
    if (k < 0 || n_rows < 0 || k % type_qk(type) != 0) return B200_ERR_SHAPE;
    if (k == 0 || n_rows == 0) return B200_OK;
    cudaStream_t st = (cudaStream_t) stream;
    unsigned * bad = b200_error_word_dev();
    // gridDim.y is limited to 65535: walk the output rows in slabs
    for (int64_t r0 = 0; r0 < n_rows; r0 += 65535) {
        const int64_t nr = n_rows - r0 < 65535 ? n_rows - r0 : 65535;
        const dim3 grid((unsigned) ceil_div(k / 8, 256), (unsigned) nr);
        const uint8_t * rows = (const uint8_t *) rows_dev + r0 * rows_stride_bytes;
        uint8_t * dst = (uint8_t *) dst_dev + r0 * dst_row_stride_bytes;
#define CALL(T, OUT) k_get_rows<T, OUT><<<grid, 256, 0, st>>>((const uint8_t *) table_dev, (int64_t) table_row_stride_bytes, k, table_rows, \
                                                             rows, (int64_t) rows_stride_bytes, dst, (int64_t) dst_row_stride_bytes, bad);
        DISPATCH_TYPE_OUT(type, out_type, CALL)
#undef CALL
        b200_count_launch(1);
    }
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}
