// mmq_tc.cuh — batched quantized MUL_MAT on the 5th-gen tensor cores (tcgen05 + TMEM).
//
//   dst[m, n] = W[m, k] (ggml quant blocks)  x  X[n, k]^T (f32)        for n > 8 (prefill)
//
// Replaces mul_mat_q / load_tiles_* / vec_dot_*_mma (reference src/ggml-cuda/mmq.cuh:277-3522,
// mmq.cu:71-143), which unpack weights to int8 and issue one mma.sync m16n8k32 per 32-wide block,
// rescaling in registers after every block.
//
// Design.  A per-32 (or per-16) block scale cannot ride through an integer TMEM accumulator: with
// kind::i8 every 32-deep slice needs its own accumulator drained and FMA-scaled on the CUDA cores
// (that route is built and measured in mmq_i8.cuh).  Here the scale is folded into the WEIGHT operand
// instead: producer warps dequantize the quant blocks to bf16 (exact integer -> bf16 via the 0x4300
// magic, one HFMA2.BF16 per two weights) straight into the UMMA canonical K-major SWIZZLE_128B
// shared-memory layout, activations are converted once to bf16 and fetched by TMA with the same
// swizzle, and a single tcgen05.mma.kind::f16 chain accumulates the whole K extent in f32 in TMEM.
// bf16, not fp16: both operands then have the exponent range of f32, so the result is as indifferent
// to the magnitude of activations and block scales (|x| ~ 1e5 after un-normalised residual sums, d * sc
// up to 65504 * 63, subnormal d) as the reference's per-block q8_1 route is (src/ggml-cuda/quantize.cu:
// 50-143); fp16 operands overflowed to inf / flushed to zero there.  8 significand bits per operand is
// the precision class of q8_1 itself (<= 7 bits per 32-block): measured NMSE ~5e-6 against 5e-4 allowed.
// kind::f16 rejects mixed f16 x bf16 operands (measured: illegal instruction), so both are bf16.
//
// CTA = one 128 x BN output tile, 10 warps:
//   warps 0-7  weight producers (two threads per row: each expands 32 of every 64 elements), then the
//              epilogue (TMEM lane quarter = warp id % 4, the pair splits the columns)
//   warp  8    activation loader (TMA 2-D tiles, zero-filled past n and k)
//   warp  9    TMEM allocation + the single MMA-issuing thread
// Raw quant bytes are prefetched two 256-deep slabs ahead by 128 per-row bulk async copies
// (cp.async.bulk, one per producer thread), re-aligned in shared memory when rows are not 16-byte
// aligned (Q6_K, odd block counts), then expanded 64 K-elements at a time into a ring of A stages.
#pragma once
#include "common.cuh"
#include "mmvq_stream.cuh"   // mbarrier / bulk-copy PTX helpers

#include <cuda.h>

namespace b200 {

// ---------------------------------------------------------------- tcgen05 PTX
__device__ __forceinline__ void tmem_alloc(uint32_t * dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after()  { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem], 128 x N x 16, f16/bf16 inputs, f32 accumulate
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on an mbarrier when all previously issued MMAs of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t * bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void mbar_arrive(uint64_t * bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
// TMA 2-D tile load (global -> swizzled shared), completion on an mbarrier
__device__ __forceinline__ void tma_load_2d(void * dst, const CUtensorMap * tmap, int c0, int c1, uint64_t * bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 :: "r"(smem_u32(dst)), "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}

// UMMA shared-memory descriptor, K-major, SWIZZLE_128B, 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t umma_desc_sw128(const void * smem_ptr) {
    const uint64_t addr = (uint64_t) (smem_u32(smem_ptr) >> 4) & 0x3FFFull;
    return addr | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// ---------------------------------------------------------------- exact small-int -> bf16 helpers
// (x & mask) | 0x4300 in each 16-bit lane is the bf16 value 128 + x for 0 <= x < 128 (7 stored mantissa bits)
typedef __nv_bfloat162 wt2;                  // two weights / activations as the tensor core reads them
__device__ __forceinline__ wt2 u32_as_h2(uint32_t v) { return *reinterpret_cast<wt2 *>(&v); }
__device__ __forceinline__ uint32_t h2_as_u32(wt2 v) { return *reinterpret_cast<uint32_t *>(&v); }
__device__ __forceinline__ wt2 f32_to_wt2(float f) { return __float2bfloat162_rn(f); }
// bytes (b0,b1) or (b2,b3) of w spread to the low byte of each 16-bit lane
__device__ __forceinline__ uint32_t spread01(uint32_t w) { return __byte_perm(w, 0, 0x4140); }
__device__ __forceinline__ uint32_t spread23(uint32_t w) { return __byte_perm(w, 0, 0x4342); }
// the same with the byte's sign replicated into the high byte of the lane (prmt's sign-replicate selector bit,
// which __byte_perm masks off): each lane = the byte sign-extended to 16 bits
__device__ __forceinline__ uint32_t spread01_sx(uint32_t w) { uint32_t r; asm("prmt.b32 %0, %1, %2, 0x9180;" : "=r"(r) : "r"(w), "r"(0u)); return r; }
__device__ __forceinline__ uint32_t spread23_sx(uint32_t w) { uint32_t r; asm("prmt.b32 %0, %1, %2, 0xB3A2;" : "=r"(r) : "r"(w), "r"(0u)); return r; }
__device__ __forceinline__ wt2 magic_to_h2(uint32_t lanes, float bias) {
    // lanes hold integers < 128 in each 16-bit half; returns (int - bias) exactly (|int - bias| < 256)
    return __hsub2(u32_as_h2(lanes | 0x43004300u), f32_to_wt2(128.0f + bias));
}

} // namespace b200

namespace b200 {

// ================================================================= weight expansion (quant -> fp16)
// A producer thread owns one weight row.  Its 256-element slab sits 16-byte aligned in shared
// memory (SLAB bytes, PITCH apart); expand64<T>(SUB) turns elements [64*SUB, 64*SUB+64) into eight
// 16-byte chunks of fp16 and stores them into the row of the SWIZZLE_128B A tile.
template <int T> struct slab;
template <> struct slab<T_Q4_0> { static constexpr int bytes = 144, pitch = 144, pitch_unaligned = 176; };
template <> struct slab<T_Q8_0> { static constexpr int bytes = 272, pitch = 272, pitch_unaligned = 304; };
template <> struct slab<T_Q4_K> { static constexpr int bytes = 144, pitch = 144, pitch_unaligned = 176; };
template <> struct slab<T_Q5_K> { static constexpr int bytes = 176, pitch = 176, pitch_unaligned = 208; };
template <> struct slab<T_Q6_K> { static constexpr int bytes = 210, pitch = 240, pitch_unaligned = 240; };
// 2-D TMA tile of one raw slab (rows 16-byte aligned in memory): box width in 32-bit words and first word of slab s.
// A TMA box must start on a 16-byte boundary of the row.  Q6_K blocks are 210 bytes: slab s starts (210 s) % 16 = 0, 2, ..
// 14 bytes past one; the box is 60 words (240 bytes >= 14 + 210, an odd multiple of 16: conflict-free 16-byte shared
// loads) and the expander reads its words `skew` bytes further in (word offset as a template parameter, the remaining
// 0 / 2 bytes by funnel shift).
template <int T> struct slab_tma {
    static constexpr int box_words = T == T_Q6_K ? 60 : slab<T>::bytes / 4;
    static __host__ __device__ __forceinline__ int word0(int s) { return T == T_Q6_K ? ((210 * s) >> 4) << 2 : s * (slab<T>::bytes / 4); }
    static __host__ __device__ __forceinline__ int skew(int s)  { return T == T_Q6_K ? (210 * s) & 15 : 0; }
};

// words [W0, W0+N) of the slab via 16-byte shared loads (conflict-free: PITCH/16 is odd)
// the same for data that starts `sh` bits (0 or 16) into word W0: one more word, funnel shifts
template <int W0, int N>
__device__ __forceinline__ void slab_words(const uint8_t * s, uint32_t (&w)[N]);
template <int W0, int N>
__device__ __forceinline__ void slab_words_sh(const uint8_t * s, uint32_t sh, uint32_t (&w)[N]) {
    uint32_t t[N + 1];
    slab_words<W0, N + 1>(s, t);
#pragma unroll
    for (int i = 0; i < N; i++) w[i] = __funnelshift_r(t[i], t[i + 1], sh);
}
template <int W0, int N>
__device__ __forceinline__ void slab_words(const uint8_t * s, uint32_t (&w)[N]) {
    constexpr int C0 = W0 / 4, C1 = (W0 + N + 3) / 4;
    uint32_t t[(C1 - C0) * 4];
#pragma unroll
    for (int c = C0; c < C1; c++) {
        const uint4 v = lds128(s + 16 * c);
        t[(c - C0) * 4 + 0] = v.x; t[(c - C0) * 4 + 1] = v.y; t[(c - C0) * 4 + 2] = v.z; t[(c - C0) * 4 + 3] = v.w;
    }
#pragma unroll
    for (int i = 0; i < N; i++) w[i] = t[W0 - 4 * C0 + i];
}

__device__ __forceinline__ void store_chunk(uint8_t * a_row, int row7, int c, wt2 v0, wt2 v1, wt2 v2, wt2 v3) {
    uint4 o;
    o.x = h2_as_u32(v0); o.y = h2_as_u32(v1); o.z = h2_as_u32(v2); o.w = h2_as_u32(v3);
    *reinterpret_cast<uint4 *>(a_row + ((c ^ row7) << 4)) = o;
}
__device__ __forceinline__ void store_zero_chunk(uint8_t * a_row, int row7, int c) {
    *reinterpret_cast<uint4 *>(a_row + ((c ^ row7) << 4)) = make_uint4(0, 0, 0, 0);
}

template <int T> struct expander;

// ---- Q4_K / Q5_K
template <bool HAS_QH> struct expander_k45 {
    uint32_t hdr[4];      // d|dmin, scales[12]
    uint32_t qh[8];       // Q5_K only
    __device__ __forceinline__ void begin_slab(const uint8_t * s, int /*nvalid*/) {
        slab_words<0, 4>(s, hdr);
        if (HAS_QH) slab_words<4, 8>(s, qh);
    }
    // half: which 32 of the 64 elements this thread expands (two producer threads per row)
    template <int SUB> __device__ __forceinline__ void expand64(const uint8_t * s, uint8_t * a_row, int row7, int half, int ccm) {
        const float d = f16_bits_to_f32(hdr[0]), dmin = f16_bits_to_f32(hdr[0] >> 16);
        uint32_t q[8];
        slab_words<(HAS_QH ? 12 : 4) + 8 * SUB, 8>(s, q);
#pragma unroll
        for (int hi = 0; hi < 2; hi++) {
            if (hi != half) continue;
            int sc, mn;
            k4_scale_min(hdr[1], hdr[2], hdr[3], 2 * SUB + hi, sc, mn);
            const wt2 d1 = f32_to_wt2(d * (float) sc);
            const wt2 m1 = f32_to_wt2(-(dmin * (float) mn));
#pragma unroll
            for (int cc = 0; cc < 4; cc++) {
                if (!((ccm >> cc) & 1)) continue;
                wt2 v[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint32_t w  = q[2 * cc + (j >> 1)];
                    uint32_t lanes = (((j & 1) ? spread23(w) : spread01(w)) >> (4 * hi)) & 0x000F000Fu;
                    if (HAS_QH) {
                        const uint32_t h = qh[2 * cc + (j >> 1)];
                        lanes |= ((((j & 1) ? spread23(h) : spread01(h)) >> (2 * SUB + hi)) & 0x00010001u) << 4;
                    }
                    v[j] = __hfma2(magic_to_h2(lanes, 0.0f), d1, m1);
                }
                store_chunk(a_row, row7, 4 * hi + cc, v[0], v[1], v[2], v[3]);
            }
        }
    }
};
template <> struct expander<T_Q4_K> : expander_k45<false> {};
template <> struct expander<T_Q5_K> : expander_k45<true> {};

// ---- Q6_K
template <> struct expander<T_Q6_K> {
    uint32_t sc[4];       // 16 int8 scales
    float d;
    int wo = 0;           // the slab starts 4 wo + sh / 8 bytes into its slot (2-D TMA tiles of 210-byte blocks: slab_tma<>)
    uint32_t sh = 0;
    __device__ __forceinline__ void set_skew(int sk) { wo = sk >> 2; sh = (uint32_t) (sk & 3) * 8u; }
    template <int WO> __device__ __forceinline__ void begin_slab_wo(const uint8_t * s) {
        uint32_t t[5];
        slab_words_sh<48 + WO, 5>(s, sh, t);        // bytes 192..211: scales[16], d
        sc[0] = t[0]; sc[1] = t[1]; sc[2] = t[2]; sc[3] = t[3];
        d = f16_bits_to_f32(t[4]);
    }
    __device__ __forceinline__ void begin_slab(const uint8_t * s, int /*nvalid*/) {
        switch (wo) {                               // uniform over the CTA
            case 0: begin_slab_wo<0>(s); break;
            case 1: begin_slab_wo<1>(s); break;
            case 2: begin_slab_wo<2>(s); break;
            default: begin_slab_wo<3>(s); break;
        }
    }
    template <int SUB> __device__ __forceinline__ void expand64(const uint8_t * s, uint8_t * a_row, int row7, int half, int ccm) {
        switch (wo) {
            case 0: expand64_wo<SUB, 0>(s, a_row, row7, half, ccm); break;
            case 1: expand64_wo<SUB, 1>(s, a_row, row7, half, ccm); break;
            case 2: expand64_wo<SUB, 2>(s, a_row, row7, half, ccm); break;
            default: expand64_wo<SUB, 3>(s, a_row, row7, half, ccm); break;
        }
    }
    template <int SUB, int WO> __device__ __forceinline__ void expand64_wo(const uint8_t * s, uint8_t * a_row, int row7, int half, int ccm) {
        constexpr int h = SUB >> 1, p = SUB & 1;
        uint32_t ql[16], qh[8];
        slab_words_sh<16 * h + WO, 16>(s, sh, ql);           // ql[64h .. 64h+64)
        slab_words_sh<32 + 8 * h + WO, 8>(s, sh, qh);        // qh[32h .. 32h+32)
#pragma unroll
        for (int c = 0; c < 8; c++) {
            if ((c >> 2) != half || !((ccm >> (c & 3)) & 1)) continue;
            const int st = 2 * p + (c >> 2);      // stream
            const int l0 = 8 * (c & 3);
            const int sidx = 8 * h + 2 * st + (l0 >> 4);
            const int scv = (int) (int8_t) ((sc[sidx >> 2] >> (8 * (sidx & 3))) & 0xFFu);
            const wt2 ds = f32_to_wt2(d * (float) scv);
            wt2 v[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint32_t wl = ql[8 * (c >> 2) + 2 * (c & 3) + (j >> 1)];
                const uint32_t wh = qh[2 * (c & 3) + (j >> 1)];
                const uint32_t lo = (((j & 1) ? spread23(wl) : spread01(wl)) >> (4 * p)) & 0x000F000Fu;
                const uint32_t hi = ((((j & 1) ? spread23(wh) : spread01(wh)) >> (2 * st)) & 0x00030003u) << 4;
                v[j] = __hmul2(magic_to_h2(lo | hi, 32.0f), ds);
            }
            store_chunk(a_row, row7, c, v[0], v[1], v[2], v[3]);
        }
    }
};

// ---- Q4_0: two 18-byte blocks per 64 elements, words 9*SUB .. 9*SUB+8 of the slab
template <> struct expander<T_Q4_0> {
    int nvalid;   // valid blocks in this slab (k tail)
    __device__ __forceinline__ void begin_slab(const uint8_t *, int nv) { nvalid = nv; }
    template <int SUB> __device__ __forceinline__ void expand64(const uint8_t * s, uint8_t * a_row, int row7, int half, int ccm) {
        uint32_t w[9];
        slab_words<9 * SUB, 9>(s, w);
#pragma unroll
        for (int b = 0; b < 2; b++) {
            if (b != half) continue;
            if (2 * SUB + b >= nvalid) {
#pragma unroll
                for (int cc = 0; cc < 4; cc++) if ((ccm >> cc) & 1) store_zero_chunk(a_row, row7, 4 * b + cc);
                continue;
            }
            uint32_t qs[4];
#pragma unroll
            for (int i = 0; i < 4; i++) qs[i] = b == 0 ? __funnelshift_r(w[i], w[i + 1], 16) : w[5 + i];
            const wt2 d = f32_to_wt2(f16_bits_to_f32(b == 0 ? w[0] : (w[4] >> 16)));
#pragma unroll
            for (int cc = 0; cc < 4; cc++) {
                if (!((ccm >> cc) & 1)) continue;
                wt2 v[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint32_t x = qs[2 * (cc & 1) + (j >> 1)];
                    const uint32_t lanes = (((j & 1) ? spread23(x) : spread01(x)) >> (4 * (cc >> 1))) & 0x000F000Fu;
                    v[j] = __hmul2(magic_to_h2(lanes, 8.0f), d);
                }
                store_chunk(a_row, row7, 4 * b + cc, v[0], v[1], v[2], v[3]);
            }
        }
    }
};

// ---- Q8_0: two 34-byte blocks per 64 elements, words 17*SUB .. 17*SUB+16
template <> struct expander<T_Q8_0> {
    int nvalid;
    __device__ __forceinline__ void begin_slab(const uint8_t *, int nv) { nvalid = nv; }
    template <int SUB> __device__ __forceinline__ void expand64(const uint8_t * s, uint8_t * a_row, int row7, int half, int ccm) {
        uint32_t w[17];
        slab_words<17 * SUB, 17>(s, w);
#pragma unroll
        for (int b = 0; b < 2; b++) {
            if (b != half) continue;
            if (2 * SUB + b >= nvalid) {
#pragma unroll
                for (int cc = 0; cc < 4; cc++) if ((ccm >> cc) & 1) store_zero_chunk(a_row, row7, 4 * b + cc);
                continue;
            }
            // int8 q = (q & 127) - 128 * sign: bf16 holds 128 + (q & 127) exactly (the magic), the sign bit of the
            // sign-extended lane, masked at bit 14, IS bf16 2.0 -> t2 = -128 d (1 + sign), v = (128 + (q & 127)) d + t2:
            // every term exact, one rounding
            const wt2 d = f32_to_wt2(f16_bits_to_f32(b == 0 ? w[0] : (w[8] >> 16)));
            const wt2 dm64 = __hmul2(d, f32_to_wt2(-64.0f)), dm128 = __hmul2(d, f32_to_wt2(-128.0f));
#pragma unroll
            for (int cc = 0; cc < 4; cc++) {
                if (!((ccm >> cc) & 1)) continue;
                wt2 v[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int wi = 2 * cc + (j >> 1);
                    const uint32_t x = b == 0 ? __funnelshift_r(w[wi], w[wi + 1], 16) : w[9 + wi];
                    const uint32_t sx = (j & 1) ? spread23_sx(x) : spread01_sx(x);
                    const wt2 t2 = __hfma2(u32_as_h2(sx & 0x40004000u), dm64, dm128);
                    v[j] = __hfma2(u32_as_h2((sx & 0x007F007Fu) | 0x43004300u), d, t2);
                }
                store_chunk(a_row, row7, 4 * b + cc, v[0], v[1], v[2], v[3]);
            }
        }
    }
};

} // namespace b200

namespace b200 {

// ================================================================= the kernel
struct mmq_args {
    const uint8_t * w;         // weight rows
    int64_t w_row_stride;      // bytes
    int64_t m, k, n;
    float * dst;               // dst[j * dst_col_stride + i]
    int64_t dst_col_stride;    // floats
    int     nblk;              // blocks per row (k / qk)
    int     raw_pitch;         // bytes per row slot in the raw ring
    // split K (matrices with fewer tiles than SMs): CTA z of a tile contracts 256-deep slabs [z * slabs_per_split, ..) and
    // writes its partial tile to part[z][j][i] (column stride m); k_mmq_reduce adds the partials in split order
    int     ksplit, slabs_per_split;
    float * part;
};

constexpr int MMQ_BM = 128;
constexpr int MMQ_BK = 64;           // K elements per MMA stage (one 128-byte swizzle row)
constexpr int MMQ_RS = 3;            // raw slab ring depth (2 for the formats with slabs > 200 bytes: shared memory)
template <int T, int BN = 256> struct mmq_rs { static constexpr int value = (slab<T>::bytes > 200 || BN > 256) ? 2 : MMQ_RS; };
#ifndef B200_MMQ_NP
#define B200_MMQ_NP 2
#endif
constexpr int MMQ_NP = B200_MMQ_NP;   // producer threads per weight row: 2 (each expands 32 of every 64 elements) or 4 (16 each:
                                      // -DB200_MMQ_NP=4, measured 5 % slower at 256-column tiles, profiles/r03u_prefill_16_producer_warps.txt)
constexpr int MMQ_PW = 4 * MMQ_NP;    // weight-producer warps
constexpr int MMQ_THREADS = (MMQ_PW + 2) * 32;      // + activation loader + MMA issuer
#ifndef MMQ_STAGES_SMALL
#define MMQ_STAGES_SMALL 4
#endif
#ifndef MMQ_CTAS_SMALL
#define MMQ_CTAS_SMALL 1
#endif

template <int BN> struct mmq_cfg {
    // (tried: 2 stages + 2 CTAs per SM for BN <= 128 to double the weight-expansion producers per tensor pipe:
    //  no gain at n = 512, profiles/r01g_prefill_probe.txt; -DMMQ_STAGES_SMALL=2 -DMMQ_CTAS_SMALL=2 rebuilds it)
    // BN = 512: two 256-column accumulators (all 512 TMEM columns) fed by ONE expansion of the weights -- the weight
    // producers, not the tensor pipe, bound this kernel (~1200 cycles per 64-deep stage whatever BN), so the second
    // accumulator's MMAs are nearly free; its 64 KB activation stages leave room for two stages only
    static constexpr int stages   = BN <= 128 ? MMQ_STAGES_SMALL : (BN > 256 ? 2 : 3);
    static constexpr int ctas_per_sm = BN <= 128 ? MMQ_CTAS_SMALL : 1;
    static constexpr int a_bytes  = MMQ_BM * 128;
    static constexpr int b_bytes  = BN * 128;
    static constexpr int tmem_cols = BN < 32 ? 32 : BN;
    static constexpr int halves   = BN > 256 ? 2 : 1;          // MMAs of N = BN / halves (N <= 256 per instruction)
};

// dynamic shared memory map (offsets from a 1024-byte aligned base)
template <int BN> struct mmq_smem {
    static constexpr int a_off    = 0;
    static constexpr int b_off    = a_off + mmq_cfg<BN>::stages * mmq_cfg<BN>::a_bytes;
    static constexpr int raw_off  = b_off + mmq_cfg<BN>::stages * mmq_cfg<BN>::b_bytes;
    static size_t total(int raw_pitch, int rs) { return (size_t) raw_off + (size_t) rs * MMQ_BM * raw_pitch + 256 + 1024; }
};

template <int T, int BN, bool ALIGNED>
__global__ void __launch_bounds__(MMQ_THREADS, mmq_cfg<BN>::ctas_per_sm)
k_mmq_tc(const mmq_args a, const __grid_constant__ CUtensorMap tmap_b, const __grid_constant__ CUtensorMap tmap_w) {
    using CFG = mmq_cfg<BN>;
    using SM  = mmq_smem<BN>;
    constexpr int ST = CFG::stages;
    constexpr int SLAB_ELEMS = 256;
    constexpr int BLK_PER_SLAB = SLAB_ELEMS / fmt<T>::qk;       // 1 or 8
    constexpr int RS = mmq_rs<T, BN>::value;
    constexpr int NH = CFG::halves, BNH = BN / NH;

    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    uint8_t * a_tiles = smem + SM::a_off;
    uint8_t * b_tiles = smem + SM::b_off;
    uint8_t * raw     = smem + SM::raw_off;
    uint8_t * ctl     = raw + (size_t) RS * MMQ_BM * a.raw_pitch;
    uint64_t * full      = reinterpret_cast<uint64_t *>(ctl);            // [ST]   A written + B landed
    uint64_t * empty     = full + ST;                                    // [ST]   MMAs of the stage retired
    uint64_t * raw_full  = empty + ST;                                   // [RS][4] raw slab landed, per producer warp
    uint64_t * raw_empty = raw_full + MMQ_RS * 4;                        // [RS]   every producer is done with the raw slab (ALIGNED)
    uint64_t * tmem_full = raw_empty + MMQ_RS;                           // accumulator complete
    uint32_t * tmem_slot = reinterpret_cast<uint32_t *>(tmem_full + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t m0 = (int64_t) blockIdx.x * MMQ_BM;
    const int64_t n0 = (int64_t) blockIdx.y * BN;
    // this CTA's share of K: slabs [s0, s0 + nslab), stages [kt0, kt0 + nkt) -- ring slots and parities count from 0
    const int nslab_all = (int) ((a.k + SLAB_ELEMS - 1) / SLAB_ELEMS);
    const int s0    = a.ksplit > 1 ? (int) blockIdx.z * a.slabs_per_split : 0;
    const int nslab = a.ksplit > 1 ? min(a.slabs_per_split, nslab_all - s0) : nslab_all;
    const int kt0   = 4 * s0;
    const int nkt   = min((int) ((a.k + MMQ_BK - 1) / MMQ_BK) - kt0, 4 * nslab);      // 64-deep MMA stages

    if (tid == 0) {
        for (int s = 0; s < ST; s++) { mbar_init(&full[s], 128 * MMQ_NP + 1); mbar_init(&empty[s], 1); }
        // ALIGNED: one 2-D TMA tile per raw slab (128 rows x slab bytes), issued by the loader thread;
        // otherwise one bulk copy per row, issued by that row's producer thread
        for (int s = 0; s < MMQ_RS * 4; s++) mbar_init(&raw_full[s], ALIGNED ? 1 : 32);
        for (int s = 0; s < MMQ_RS; s++) mbar_init(&raw_empty[s], 128 * MMQ_NP);
        mbar_init(tmem_full, 1);
        mbar_fence_init();
    }
    if (warp == MMQ_PW + 1) tmem_alloc(tmem_slot, CFG::tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < MMQ_PW) {
        // ===================================================== weight producers: thread pair (prow, half) per row.
        // The quant -> fp16 expansion is the bound of this kernel (one warp per scheduler cannot hide its own
        // latencies); warps w and w + 4 share rows 32 w .. 32 w + 31, each expanding 32 of every 64 elements.
        // thread (prow, part): part = which 64 / MMQ_NP elements of every 64 it expands; the part-0 thread of a row also
        // issues / realigns its raw bytes (unaligned rows) 
        const int prow = tid & 127, part = tid >> 7, pw = warp & 3;
        const int half = part / (MMQ_NP / 2), ccm = MMQ_NP == 4 ? ((part & 1) ? 0xC : 0x3) : 0xF;
        const int64_t row = (m0 + prow < a.m) ? (m0 + prow) : (a.m - 1);   // rows past m recompute the last row, never stored
        const uint8_t * wrow = a.w + row * a.w_row_stride;
        const int row7 = prow & 7;
        uint8_t * my_raw[RS];
#pragma unroll
        for (int r = 0; r < RS; r++) my_raw[r] = raw + ((size_t) r * MMQ_BM + prow) * a.raw_pitch;

        auto slab_valid_blocks = [&](int s) {            // s: index within this CTA's share
            const int left = a.nblk - (s0 + s) * BLK_PER_SLAB;
            return left < BLK_PER_SLAB ? left : BLK_PER_SLAB;
        };
        auto issue_raw = [&](int s) {
            const int r = s % RS;
            const uint8_t * src = wrow + (size_t) (s0 + s) * BLK_PER_SLAB * fmt<T>::bytes;
            const uint32_t nbytes = (uint32_t) slab_valid_blocks(s) * fmt<T>::bytes;
            const uint32_t off = ALIGNED ? 0u : (uint32_t) (reinterpret_cast<uintptr_t>(src) & 15);
            const uint32_t cp = (off + nbytes + 15u) & ~15u;
            uint64_t * bar = &raw_full[r * 4 + pw];
            mbar_arrive_expect_tx(bar, cp);
            bulk_g2s(my_raw[r], src - off, cp, bar);
        };

        if (!ALIGNED && part == 0) {
            for (int s = 0; s < RS && s < nslab; s++) issue_raw(s);
        }

        expander<T> ex;
        for (int s = 0; s < nslab; s++) {
            const int r = s % RS;
            mbar_wait(&raw_full[r * 4 + (ALIGNED ? 0 : pw)], (uint32_t) ((s / RS) & 1));
            uint8_t * slab_ptr = my_raw[r];
            if (!ALIGNED && part == 0) {
                // shift every row of this warp down to its 16-byte aligned slot start (2-byte granules,
                // lanes side by side: conflict-free), in place
                const uint8_t * src = wrow + (size_t) (s0 + s) * BLK_PER_SLAB * fmt<T>::bytes;
                const int my_off = (int) (reinterpret_cast<uintptr_t>(src) & 15);
                constexpr int NU16 = (slab<T>::bytes + 1) / 2;
                for (int rr = 0; rr < 32; rr++) {
                    const int off = __shfl_sync(0xFFFFFFFFu, my_off, rr);
                    if (off == 0) continue;
                    uint8_t * rp = raw + ((size_t) r * MMQ_BM + pw * 32 + rr) * a.raw_pitch;
#pragma unroll
                    for (int i0 = 0; i0 < NU16; i0 += 32) {
                        const int i = i0 + lane;
                        uint16_t v = 0;
                        if (i < NU16) v = *reinterpret_cast<const uint16_t *>(rp + off + 2 * i);
                        __syncwarp();
                        if (i < NU16) *reinterpret_cast<uint16_t *>(rp + 2 * i) = v;
                        __syncwarp();
                    }
                }
                __syncwarp();
            }
            if (!ALIGNED) named_bar_sync(1 + pw, 32 * MMQ_NP);        // the pair's other warp reads the realigned rows
            if constexpr (T == T_Q6_K) ex.set_skew(ALIGNED ? slab_tma<T>::skew(s0 + s) : 0);
            ex.begin_slab(slab_ptr, slab_valid_blocks(s));
#pragma unroll
            for (int sub = 0; sub < 4; sub++) {
                const int kt = s * 4 + sub;
                if (kt < nkt) {
                    const int st = kt % ST;
                    if (kt >= ST) mbar_wait(&empty[st], (uint32_t) (((kt / ST) - 1) & 1));
                    uint8_t * a_row = a_tiles + (size_t) st * CFG::a_bytes + prow * 128;
                    if      (sub == 0) ex.template expand64<0>(slab_ptr, a_row, row7, half, ccm);
                    else if (sub == 1) ex.template expand64<1>(slab_ptr, a_row, row7, half, ccm);
                    else if (sub == 2) ex.template expand64<2>(slab_ptr, a_row, row7, half, ccm);
                    else               ex.template expand64<3>(slab_ptr, a_row, row7, half, ccm);
                    fence_proxy_async_smem();       // generic-proxy stores -> visible to the tensor core's async proxy
                    mbar_arrive(&full[st]);
                }
            }
            if (ALIGNED) {
                mbar_arrive(&raw_empty[r]);          // the loader refills the slot when all 256 producers have left it
            } else {
                named_bar_sync(1 + pw, 32 * MMQ_NP);          // both warps of the pair are done with the raw slot
                if (part == 0 && s + RS < nslab) issue_raw(s + RS);
            }
        }

        // ===================================================== epilogue: TMEM -> registers -> global
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        const int64_t out_row = m0 + prow;
        float * out = a.ksplit > 1 ? a.part + (size_t) blockIdx.z * (size_t) a.n * (size_t) a.m : a.dst;
        const int64_t out_stride = a.ksplit > 1 ? a.m : a.dst_col_stride;
        const uint32_t taddr = tmem_base + ((uint32_t) (pw * 32) << 16);      // a warp reaches TMEM lanes 32 (warp % 4) ..
        constexpr int CH = BN / MMQ_NP >= 8 ? BN / MMQ_NP : 8;                  // .. the row's MMQ_NP threads split the columns
#pragma unroll 1
        for (int c0 = part * CH; c0 < (part + 1) * CH && c0 < BN; c0 += 8) {
            if (n0 + c0 >= a.n) break;               // warp-uniform
            uint32_t v[8];
            tmem_ld8(taddr + (uint32_t) c0, v);
            tmem_ld_wait();
            if (out_row < a.m) {
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    if (n0 + c0 + j < a.n) out[(n0 + c0 + j) * out_stride + out_row] = __uint_as_float(v[j]);
                }
            }
        }
        tc_fence_before();
    } else if (warp == MMQ_PW) {
        // ===================================================== activation loader (TMA)
        if (lane == 0) {
            int next_slab = 0;
            for (int kt = 0; kt < nkt; kt++) {
                if (ALIGNED) {
                    // raw weight slabs, MMQ_RS - 1 ahead of the slab the producers expand: one 2-D TMA tile each (rows
                    // past m and words past the row end are zero-filled)
                    while (next_slab < nslab && next_slab <= kt / 4 + RS - 1) {
                        const int r = next_slab % RS;
                        if (next_slab >= RS) mbar_wait(&raw_empty[r], (uint32_t) (((next_slab / RS) - 1) & 1));
                        mbar_arrive_expect_tx(&raw_full[r * 4], (uint32_t) (MMQ_BM * slab_tma<T>::box_words * 4));
                        tma_load_2d(raw + (size_t) r * MMQ_BM * a.raw_pitch, &tmap_w, slab_tma<T>::word0(s0 + next_slab), (int) m0, &raw_full[r * 4]);
                        next_slab++;
                    }
                }
                const int st = kt % ST;
                if (kt >= ST) mbar_wait(&empty[st], (uint32_t) (((kt / ST) - 1) & 1));
                mbar_arrive_expect_tx(&full[st], (uint32_t) CFG::b_bytes);
#pragma unroll
                for (int h = 0; h < NH; h++)       // a TMA box has at most 256 rows
                    tma_load_2d(b_tiles + (size_t) st * CFG::b_bytes + (size_t) h * BNH * 128, &tmap_b, (kt0 + kt) * MMQ_BK, (int) n0 + h * BNH, &full[st]);
            }
        }
    } else {
        // ===================================================== MMA issuer
        if (lane == 0) {
            // instruction descriptor: D=f32, A=bf16, B=bf16 (kind::f16 rejects mixed f16 x bf16 operands: measured,
            // illegal instruction), both K-major, N=BN, M=128
            constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t) (BNH >> 3) << 17) | ((uint32_t) (MMQ_BM >> 4) << 24);
            for (int kt = 0; kt < nkt; kt++) {
                const int st = kt % ST;
                mbar_wait(&full[st], (uint32_t) ((kt / ST) & 1));
                tc_fence_after();
                const uint64_t da = umma_desc_sw128(a_tiles + (size_t) st * CFG::a_bytes);
#pragma unroll
                for (int h = 0; h < NH; h++) {
                    const uint64_t db = umma_desc_sw128(b_tiles + (size_t) st * CFG::b_bytes + (size_t) h * BNH * 128);
#pragma unroll
                    for (int k4 = 0; k4 < MMQ_BK / 16; k4++) {
                        // +32 bytes (2 x 16-byte units) along K inside the 128-byte swizzle row
                        umma_f16(tmem_base + (uint32_t) (h * BNH), da + (uint64_t) (2 * k4), db + (uint64_t) (2 * k4), idesc, (kt > 0 || k4 > 0) ? 1u : 0u);
                    }
                }
                umma_commit(&empty[st]);             // frees the stage when these MMAs retire
            }
            umma_commit(tmem_full);
        }
    }

    __syncthreads();
    if (warp == MMQ_PW + 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, CFG::tmem_cols);
    }
}

// f32 [n, k] (row stride in bytes) -> bf16 [n, k] packed, RN (the full exponent range of f32 survives)
static __global__ void __launch_bounds__(256) k_f32_to_bf16(const uint8_t * __restrict__ x, int64_t x_row_stride, int64_t k, uint16_t * __restrict__ y) {
    const int64_t row = blockIdx.y;
    const int64_t i = ((int64_t) blockIdx.x * 256 + threadIdx.x) * 4;
    if (i >= k) return;
    const float * xr = reinterpret_cast<const float *>(x + row * x_row_stride);
    uint16_t * yr = y + row * k;
    auto cvt = [](float f) -> uint16_t {
        return __bfloat16_as_ushort(__float2bfloat16_rn(f));
    };
    if (i + 3 < k && (reinterpret_cast<uintptr_t>(xr + i) & 15) == 0) {
        const float4 v = *reinterpret_cast<const float4 *>(xr + i);
        uint2 o;
        o.x = (uint32_t) cvt(v.x) | ((uint32_t) cvt(v.y) << 16);
        o.y = (uint32_t) cvt(v.z) | ((uint32_t) cvt(v.w) << 16);
        *reinterpret_cast<uint2 *>(yr + i) = o;
    } else {
        for (int j = 0; j < 4 && i + j < k; j++) yr[i + j] = cvt(xr[i + j]);
    }
}

// split K: dst[j][i] = part[0][j][i] + part[1][j][i] + ..  (fixed order: the result does not depend on the grid)
static __global__ void __launch_bounds__(256) k_mmq_reduce(const float * __restrict__ part, int ksplit, int64_t m, int64_t n,
                                                           float * __restrict__ dst, int64_t dst_col_stride) {
    const int64_t j = blockIdx.y;
    const int64_t i = ((int64_t) blockIdx.x * 256 + threadIdx.x) * 4;
    if (i >= m) return;
    const float * p = part + j * m + i;
    float * d = dst + j * dst_col_stride + i;
    if (i + 3 < m && ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(d)) & 15) == 0 && ((m * n) & 3) == 0) {
        float4 acc = *reinterpret_cast<const float4 *>(p);
        for (int z = 1; z < ksplit; z++) {
            const float4 v = *reinterpret_cast<const float4 *>(p + (size_t) z * m * n);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        *reinterpret_cast<float4 *>(d) = acc;
    } else {
        for (int e = 0; e < 4 && i + e < m; e++) {
            float acc = p[e];
            for (int z = 1; z < ksplit; z++) acc += p[e + (size_t) z * m * n];
            d[e] = acc;
        }
    }
}

} // namespace b200
