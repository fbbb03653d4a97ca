// unpack.cuh — one 32-weight sub-block of any hot-path format -> 8 words of 4 x int8 ready for
// __dp4a, plus the float factors that turn the integer dot products into the real one:
//
//     sum_i w_i * x_i  =  d8 * ( sA * dot(q[0..3], y[0..3]) + sB * dot(q[4..7], y[4..7]) - mn * sum(y) )
//
// where x_i = d8 * y_i is the q8_1 activation.  Element order inside the sub-block is the natural
// one (word j holds elements 4j..4j+3), so y is just the 32 consecutive int8 activations.
//
// What this replaces in the reference: the per-format vec_dot_<fmt>_q8_1 primitives and their
// nibble / scale unpacking (src/ggml-cuda/vecdotq.cuh:59-78, 187-199, 417-439, 473-502, 536-556,
// 587-608, 669-689, 741-867) and the gfx906 overrides of the same functions
// ({q4_0,q8_0,q5_k,q6_k}-gfx906.cuh).  Unlike those, the "-8" / "-32" offsets are applied to the
// weights in integer form (no fp16 `s` term), so the result equals the exact dot product of the
// dequantized weights with the q8_1-rounded activations up to f32 summation order.
#pragma once
#include "common.cuh"

namespace b200 {

struct sub32 {
    int   q[8];   // 32 int8 weights (before scaling)
    float sA;     // scale of q[0..3]  (elements 0..15)
    float sB;     // scale of q[4..7]  (elements 16..31)
    float mn;     // offset: value = scale*q - mn
};

// Loader policy: 2-byte aligned reads from global memory.
struct ld_global_a2 {
    static __device__ __forceinline__ uint32_t u8 (const uint8_t * p) { return (uint32_t) __ldg(p); }
    static __device__ __forceinline__ uint32_t u16(const uint8_t * p) { return ld_u16(p); }
    static __device__ __forceinline__ uint32_t u32(const uint8_t * p) { return ld_u32_a2(p); }
};
// Loader policy: 2-byte aligned reads from shared (or any generic) memory.
struct ld_generic_a2 {
    static __device__ __forceinline__ uint32_t u8 (const uint8_t * p) { return (uint32_t) *p; }
    static __device__ __forceinline__ uint32_t u16(const uint8_t * p) { return (uint32_t) *reinterpret_cast<const uint16_t *>(p); }
    static __device__ __forceinline__ uint32_t u32(const uint8_t * p) {
        return u16(p) | (u16(p + 2) << 16);
    }
};

// number of 32-weight sub-blocks per block
template <int T> struct subs { static constexpr int n = fmt<T>::qk / 32; };

template <int T, typename LD> struct unpack32;

template <typename LD> struct unpack32<T_Q4_0, LD> {
    static __device__ __forceinline__ void run(const uint8_t * b, int /*s*/, sub32 & o) {
        const float d = f16_bits_to_f32(LD::u16(b));
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t w = LD::u32(b + 2 + 4 * j);
            o.q[j]     = __vsubss4((int) (w & 0x0F0F0F0Fu), 0x08080808);
            o.q[j + 4] = __vsubss4((int) ((w >> 4) & 0x0F0F0F0Fu), 0x08080808);
        }
        o.sA = d; o.sB = d; o.mn = 0.0f;
    }
};

template <typename LD> struct unpack32<T_Q8_0, LD> {
    static __device__ __forceinline__ void run(const uint8_t * b, int /*s*/, sub32 & o) {
        const float d = f16_bits_to_f32(LD::u16(b));
#pragma unroll
        for (int j = 0; j < 8; j++) o.q[j] = (int) LD::u32(b + 2 + 4 * j);
        o.sA = d; o.sB = d; o.mn = 0.0f;
    }
};

// ---- the formats only the general kernels serve: value = scale * q - mn with q >= 0 for the *_1 formats (mn = -m)
template <bool HAS_M, bool HAS_QH, bool NL, typename LD>
__device__ __forceinline__ void unpack32_x(const uint8_t * b, sub32 & o) {
    const float d = f16_bits_to_f32(LD::u16(b));
    constexpr int oqh = HAS_M ? 4 : 2, oqs = oqh + (HAS_QH ? 4 : 0);
    const uint32_t qh = HAS_QH ? LD::u32(b + oqh) : 0u;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const uint32_t w = LD::u32(b + oqs + 4 * j);
        uint32_t lo = w & 0x0F0F0F0Fu, hi = (w >> 4) & 0x0F0F0F0Fu;
        if (HAS_QH) {
            // bit 4j + i of qh -> bit 4 of byte i (elements 4j..4j+3), bit 16 + 4j + i likewise for the high nibbles
            const uint32_t bl = (qh >> (4 * j)) & 0xFu, bh = (qh >> (16 + 4 * j)) & 0xFu;
            lo |= ((bl & 1u) << 4) | ((bl & 2u) << 11) | ((bl & 4u) << 18) | ((bl & 8u) << 25);
            hi |= ((bh & 1u) << 4) | ((bh & 2u) << 11) | ((bh & 4u) << 18) | ((bh & 8u) << 25);
        }
        if (NL) {
            uint32_t tl = 0, th = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                tl |= (uint32_t) (iq4nl_value((int) ((lo >> (8 * i)) & 0xFu)) & 0xFF) << (8 * i);
                th |= (uint32_t) (iq4nl_value((int) ((hi >> (8 * i)) & 0xFu)) & 0xFF) << (8 * i);
            }
            lo = tl; hi = th;
        } else if (HAS_QH && !HAS_M) {
            lo = (uint32_t) __vsubss4((int) lo, 0x10101010); hi = (uint32_t) __vsubss4((int) hi, 0x10101010);     // Q5_0: q - 16
        }
        o.q[j] = (int) lo; o.q[j + 4] = (int) hi;
    }
    o.sA = d; o.sB = d; o.mn = HAS_M ? -f16_bits_to_f32(LD::u16(b + 2)) : 0.0f;
}
template <typename LD> struct unpack32<T_Q4_1, LD> { static __device__ __forceinline__ void run(const uint8_t * b, int, sub32 & o) { unpack32_x<true, false, false, LD>(b, o); } };
template <typename LD> struct unpack32<T_Q5_0, LD> { static __device__ __forceinline__ void run(const uint8_t * b, int, sub32 & o) { unpack32_x<false, true, false, LD>(b, o); } };
template <typename LD> struct unpack32<T_Q5_1, LD> { static __device__ __forceinline__ void run(const uint8_t * b, int, sub32 & o) { unpack32_x<true, true, false, LD>(b, o); } };
template <typename LD> struct unpack32<T_IQ4_NL, LD> { static __device__ __forceinline__ void run(const uint8_t * b, int, sub32 & o) { unpack32_x<false, false, true, LD>(b, o); } };

template <bool HAS_QH, typename LD>
__device__ __forceinline__ void unpack32_k45(const uint8_t * b, int s, sub32 & o) {
    const float d    = f16_bits_to_f32(LD::u16(b));
    const float dmin = f16_bits_to_f32(LD::u16(b + 2));
    const uint32_t s0 = LD::u32(b + 4), s1 = LD::u32(b + 8), s2 = LD::u32(b + 12);
    int sc, mn;
    k4_scale_min(s0, s1, s2, s, sc, mn);
    const int g = s >> 1, hi = s & 1;
    const uint8_t * qs = b + (HAS_QH ? 48 : 16) + 32 * g;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        uint32_t v = (LD::u32(qs + 4 * j) >> (4 * hi)) & 0x0F0F0F0Fu;
        if (HAS_QH) {
            const uint32_t h = LD::u32(b + 16 + 4 * j);
            v |= ((h >> s) & 0x01010101u) << 4;
        }
        o.q[j] = (int) v;
    }
    o.sA = d * (float) sc; o.sB = o.sA; o.mn = dmin * (float) mn;
}
template <typename LD> struct unpack32<T_Q4_K, LD> {
    static __device__ __forceinline__ void run(const uint8_t * b, int s, sub32 & o) { unpack32_k45<false, LD>(b, s, o); }
};
template <typename LD> struct unpack32<T_Q5_K, LD> {
    static __device__ __forceinline__ void run(const uint8_t * b, int s, sub32 & o) { unpack32_k45<true, LD>(b, s, o); }
};

template <typename LD> struct unpack32<T_Q6_K, LD> {
    static __device__ __forceinline__ void run(const uint8_t * b, int s, sub32 & o) {
        const int h = s >> 2, st = s & 3;
        const float d = f16_bits_to_f32(LD::u16(b + 208));
        const uint32_t scw = LD::u16(b + 192 + 8 * h + 2 * st);
        const int scA = (int) (int8_t) (scw & 0xFFu), scB = (int) (int8_t) (scw >> 8);
        const uint8_t * ql = b + 64 * h + 32 * (st & 1);
        const uint8_t * qh = b + 128 + 32 * h;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint32_t lo = (LD::u32(ql + 4 * j) >> (4 * (st >> 1))) & 0x0F0F0F0Fu;
            const uint32_t hi = ((LD::u32(qh + 4 * j) >> (2 * st)) & 0x03030303u) << 4;
            o.q[j] = __vsubss4((int) (lo | hi), 0x20202020);
        }
        o.sA = d * (float) scA; o.sB = d * (float) scB; o.mn = 0.0f;
    }
};

// dot of one unpacked sub-block with 32 int8 activations (8 words) and their scale d8
__device__ __forceinline__ float dot_sub32(const sub32 & w, const int (&y)[8], float d8) {
    int a = 0, b = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) { a = __dp4a(w.q[j], y[j], a); b = __dp4a(w.q[j + 4], y[j + 4], b); }
    float r = w.sA * (float) a + w.sB * (float) b;
    if (w.mn != 0.0f) {
        int sy = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) sy = __dp4a(0x01010101, y[j], sy);
        r -= w.mn * (float) sy;
    }
    return d8 * r;
}

} // namespace b200
