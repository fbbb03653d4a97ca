// common.cuh — block layouts of the ggml quant formats on the hot path and small device helpers.
//
// Layout contract (byte offsets; all multi-byte fields little-endian, blocks are only 2-byte
// aligned in general).  Restated from the format definitions, reference src/ggml-common.h:
//   Q4_0 :170-175   18 B / 32   f16 d | u8 qs[16]            w[j]=(qs[j]&15)-8, w[j+16]=(qs[j]>>4)-8, times d
//   Q8_0 :219-224   34 B / 32   f16 d | i8 qs[32]            w[j]=qs[j]*d
//   Q8_1 :226-237   36 B / 32   f16 d | f16 s | i8 qs[32]    activation scratch, s = d*sum(qs)
//   Q4_K :295-306  144 B / 256  f16 d | f16 dmin | u8 scales[12] | u8 qs[128]
//   Q5_K :312-324  176 B / 256  f16 d | f16 dmin | u8 scales[12] | u8 qh[32] | u8 qs[128]
//   Q6_K :330-336  210 B / 256  u8 ql[128] | u8 qh[64] | i8 scales[16] | f16 d
// and, served by the general kernels only (dequantize / GET_ROWS / CPY, the one-warp-per-row mat-vec):
//   Q4_1 :177-189   20 B / 32   f16 d | f16 m | u8 qs[16]             w = q * d + m
//   Q5_0 :191-196   22 B / 32   f16 d | u8 qh[4] | u8 qs[16]          w = ((q | h << 4) - 16) * d, h = bit j / j + 16 of qh
//   Q5_1 :198-211   24 B / 32   f16 d | f16 m | u8 qh[4] | u8 qs[16]  w = (q | h << 4) * d + m
//   IQ4_NL :427-431 18 B / 32   f16 d | u8 qs[16]                     w = kvalues_iq4nl[q] * d
#pragma once

#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ggml_b200.h"

extern "C" void b200_count_launch(uint64_t n);   // internal launch counter (capi.cu)
extern "C" int  b200_option(int id);             // run-time path switch (capi.cu; enum b200_option)
extern "C" unsigned * b200_error_word_dev(void); // device pointer of the sticky error word (capi.cu)

namespace b200 {

// numeric values of the reference's enum ggml_type (include/ggml.h:376-410)
enum : int {
    T_F32 = B200_TYPE_F32, T_F16 = B200_TYPE_F16, T_Q4_0 = B200_TYPE_Q4_0, T_Q8_0 = B200_TYPE_Q8_0,
    T_Q8_1 = B200_TYPE_Q8_1, T_Q4_K = B200_TYPE_Q4_K, T_Q5_K = B200_TYPE_Q5_K, T_Q6_K = B200_TYPE_Q6_K,
    T_BF16 = B200_TYPE_BF16, T_I32 = B200_TYPE_I32,
    T_Q4_1 = B200_TYPE_Q4_1, T_Q5_0 = B200_TYPE_Q5_0, T_Q5_1 = B200_TYPE_Q5_1, T_IQ4_NL = B200_TYPE_IQ4_NL,
};

template <int T> struct fmt;
template <> struct fmt<T_Q4_0> { static constexpr int qk = 32,  bytes = 18;  };
template <> struct fmt<T_Q8_0> { static constexpr int qk = 32,  bytes = 34;  };
template <> struct fmt<T_Q4_K> { static constexpr int qk = 256, bytes = 144; };
template <> struct fmt<T_Q5_K> { static constexpr int qk = 256, bytes = 176; };
template <> struct fmt<T_Q6_K> { static constexpr int qk = 256, bytes = 210; };
template <> struct fmt<T_Q4_1> { static constexpr int qk = 32,  bytes = 20;  };
template <> struct fmt<T_Q5_0> { static constexpr int qk = 32,  bytes = 22;  };
template <> struct fmt<T_Q5_1> { static constexpr int qk = 32,  bytes = 24;  };
template <> struct fmt<T_IQ4_NL> { static constexpr int qk = 32, bytes = 18; };

__host__ __device__ inline int type_qk(int t) {
    switch (t) {
        case T_Q4_0: case T_Q8_0: case T_Q8_1: case T_Q4_1: case T_Q5_0: case T_Q5_1: case T_IQ4_NL: return 32;
        case T_Q4_K: case T_Q5_K: case T_Q6_K: return 256;
        case T_F32: case T_F16: case T_BF16: case T_I32: return 1;
        default: return 0;
    }
}
__host__ __device__ inline int type_bytes(int t) {
    switch (t) {
        case T_Q4_0: return 18; case T_Q8_0: return 34; case T_Q8_1: return 36;
        case T_Q4_1: return 20; case T_Q5_0: return 22; case T_Q5_1: return 24; case T_IQ4_NL: return 18;
        case T_Q4_K: return 144; case T_Q5_K: return 176; case T_Q6_K: return 210;
        case T_F32: case T_I32: return 4; case T_F16: case T_BF16: return 2;
        default: return 0;
    }
}
// the five formats of the hot path (every kernel) ..
__host__ __device__ inline bool type_is_quant(int t) {
    return t == T_Q4_0 || t == T_Q8_0 || t == T_Q4_K || t == T_Q5_K || t == T_Q6_K;
}
// .. and those plus the formats only the general kernels serve
__host__ __device__ inline bool type_is_block(int t) {
    return type_is_quant(t) || t == T_Q4_1 || t == T_Q5_0 || t == T_Q5_1 || t == T_IQ4_NL;
}
// kvalues_iq4nl (src/ggml-common.h:1109-1112) as four little-endian words: entry q = byte (q & 3) of word q >> 2
__device__ __forceinline__ int iq4nl_value(int q) {
    const uint32_t w = q < 8 ? (q < 4 ? 0xBFAD9881u : 0xF6EADDCFu) : (q < 12 ? 0x26190D01u : 0x71594535u);
    return (int) (int8_t) ((w >> (8 * (q & 3))) & 0xFFu);
}

// ---------------------------------------------------------------- unaligned-safe global loads
// Blocks are 2-byte aligned; these go through the read-only path.
__device__ __forceinline__ uint32_t ld_u16(const uint8_t * p) {
    return (uint32_t) __ldg(reinterpret_cast<const uint16_t *>(p));
}
// 32-bit value from a 2-byte aligned address (two 16-bit loads, one PRMT-free OR)
__device__ __forceinline__ uint32_t ld_u32_a2(const uint8_t * p) {
    return ld_u16(p) | (ld_u16(p + 2) << 16);
}
__device__ __forceinline__ uint32_t ld_u32_a4(const uint8_t * p) {
    return __ldg(reinterpret_cast<const uint32_t *>(p));
}
// 8 bytes from a 2-byte aligned address with at most three loads; may read up to p + 9 (callers keep that
// inside the block), never before p
__device__ __forceinline__ void ld_2x32_a2(const uint8_t * p, uint32_t & w0, uint32_t & w1) {
    if ((reinterpret_cast<uintptr_t>(p) & 2) == 0) {
        w0 = ld_u32_a4(p); w1 = ld_u32_a4(p + 4);
    } else {
        const uint32_t a = ld_u16(p), b = ld_u32_a4(p + 2), c = ld_u32_a4(p + 6);
        w0 = a | (b << 16); w1 = (b >> 16) | (c << 16);
    }
}
__device__ __forceinline__ uint2 ld_u64(const uint8_t * p) {
    return __ldg(reinterpret_cast<const uint2 *>(p));
}
__device__ __forceinline__ uint4 ld_u128(const uint8_t * p) {
    return __ldg(reinterpret_cast<const uint4 *>(p));
}
__device__ __forceinline__ float f16_bits_to_f32(uint32_t h) {
    return __half2float(__ushort_as_half((unsigned short) (h & 0xFFFFu)));
}

// 6-bit (scale, min) pair j of a Q4_K/Q5_K super-block from the 12 packed bytes held in three
// words (w0 = bytes 0-3, w1 = 4-7, w2 = 8-11).  Same values as get_scale_min_k4
// (reference src/ggml-quants.c:703-710).
__device__ __forceinline__ uint32_t k4_byte(uint32_t w0, uint32_t w1, uint32_t w2, int i) {
    const uint32_t w = i < 4 ? w0 : (i < 8 ? w1 : w2);
    return (w >> (8 * (i & 3))) & 0xFFu;
}
__device__ __forceinline__ void k4_scale_min(uint32_t w0, uint32_t w1, uint32_t w2, int j, int & sc, int & mn) {
    if (j < 4) {
        sc = (int) (k4_byte(w0, w1, w2, j) & 63u);
        mn = (int) (k4_byte(w0, w1, w2, j + 4) & 63u);
    } else {
        const uint32_t hi = k4_byte(w0, w1, w2, j + 4);
        sc = (int) ((hi & 0x0Fu) | ((k4_byte(w0, w1, w2, j - 4) >> 6) << 4));
        mn = (int) ((hi >> 4)    | ((k4_byte(w0, w1, w2, j)     >> 6) << 4));
    }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
    return v;
}

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// internal (not exported) cross-TU helpers
int b200_quantize_acts_planes(const float * x_dev, size_t x_row_stride_bytes, int64_t k, int64_t n,
                              int8_t * q_plane, float * d_plane, cudaStream_t st);

// Launch-error plumbing for the C ABI: kernels never abort the process; the entry point returns
// the cudaError_t as a negative code.
#define B200_RETURN_IF_LAUNCH_FAILED()                         \
    do {                                                       \
        cudaError_t e__ = cudaGetLastError();                  \
        if (e__ != cudaSuccess) return -(int) e__;             \
    } while (0)

} // namespace b200
