// mmvq_fused.cuh — the B200 decode mat-vec for n = 1: ONE kernel per MUL_MAT.
//
//   dst[m] = W[m, k] (ggml quant blocks) . x[k] (f32)
//
// Replaces quantize_row_q8_1_cuda + mul_mat_vec_q (reference src/ggml-cuda/quantize.cu:145-157,
// mmvq.cu:139-343): the reference launches a quantize kernel, round-trips block_q8_1 through global
// memory and then runs one CTA per row with 2-4 byte loads.
//
// What this kernel does instead
//   * activations are quantized to q8_1 values IN the kernel (bit-exact quantize_row_q8_1_ref math,
//     src/ggml-quants.c:225-258) by the CTA's consumer warps into shared memory, then each lane pulls
//     the slice of K it owns into registers for the lifetime of the kernel;
//   * a dedicated producer warp streams weight rows through a ring of shared-memory stages with bulk
//     async copies (cp.async.bulk, SASS UBLKCP) on full/empty mbarriers — bytes in flight per SM are
//     set by the ring, not by occupancy;
//   * a lane owns 128 consecutive weights of a row ("item" = half a K-quant super-block): the 6-bit
//     scale unpack is paid once per 128 weights and the warp-level reduction once per row pair;
//   * W consecutive warps share a row when k > 4096 (partials meet in shared memory in fixed order:
//     results never depend on the grid, so a row-split GPU computes bit-identical rows);
//   * programmatic dependent launch: the kernel starts while its predecessor drains; when the caller
//     says the weights are constant (model weights), their first stages are requested before the
//     grid dependency resolves, so HBM stays busy across back-to-back mat-vecs.
#pragma once
#include "common.cuh"
#include "mmvq_stream.cuh"      // mbarrier / bulk-copy helpers, dev_caps, env_int

namespace b200 {

__device__ __forceinline__ void mbar_arrive_cta(uint64_t * bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void pdl_wait()   { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// dot of four unsigned bytes with four signed bytes
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// Tuning instrumentation (per-CTA globaltimer stamps, skip-the-math modes) exists only in -DB200_TUNING builds;
// the product kernel carries none of it.
#ifdef B200_TUNING
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t;
}
#define FUSED_STAMP(slot) do { if (a.dbg && lane == 0) a.dbg[(size_t) blockIdx.x * 8 + (slot)] = gtime(); } while (0)
#define FUSED_DBG_MODE(bit) (a.dbg_mode & (bit))
#else
#define FUSED_STAMP(slot) do { } while (0)
#define FUSED_DBG_MODE(bit) 0
#endif

constexpr int FUSED_RB = 2;          // rows a warp carries at once (one shuffle tree per pair)

// ---------------------------------------------------------------- q8 activations of one item
// 128 consecutive activations = four q8_1 blocks: int8 values, fp16-rounded d, d * sum(q).
struct acts128 {
    int   q[4][8];
    float d[4];       // d of block j (the odd blocks pre-divided by 16 where the format asks for it)
    float sd[4];      // d * sum(q) of block j                      (formats with a min / offset per 32)
    int   ys16[8];    // sum(q) of each 16-value half of block j       (Q6_K: one scale per 16, offset 32)
};

// ================================================================= per-format half-block items
template <int T> struct hitem;

// ---- Q4_K: block 144 B = d,dmin | scales[12] | qs[128]; item h of block b: sub-blocks 4h..4h+3,
// quant bytes qs[64h, 64h+64): byte 32g+l holds element 64g+l (low nibble) and 64g+32+l (high nibble).
template <> struct hitem<T_Q4_K> {
    static constexpr bool aligned16 = true;    // rows and items are 16-byte aligned: no skew inside a stage
    static constexpr bool need_ys16 = false;
    static constexpr bool hi_div16 = true;     // high nibbles are used in place (x16), folded into d
    static bool ok(int64_t k, int64_t row_stride, uintptr_t base) { return k % 256 == 0 && row_stride % 16 == 0 && base % 16 == 0; }
    struct wts { uint4 hdr; uint4 q[4]; };
    static __device__ __forceinline__ void load(wts & w, const uint8_t * row, int idx) {
        const uint8_t * bp = row + (idx >> 1) * 144;
        w.hdr = lds128(bp);
        const uint8_t * qp = bp + 16 + 64 * (idx & 1);
#pragma unroll
        for (int i = 0; i < 4; i++) w.q[i] = lds128(qp + 16 * i);
    }
    static __device__ __forceinline__ float dot(const wts & w, int h, const acts128 & a) {
        const float d = f16_bits_to_f32(w.hdr.x), dmin = f16_bits_to_f32(w.hdr.x >> 16);
        // packed 6-bit scales / mins of sub-blocks 4h..4h+3 (get_scale_min_k4, src/ggml-quants.c:703-710)
        const uint32_t s0 = w.hdr.y, s1 = w.hdr.z, s2 = w.hdr.w;
        const uint32_t sc4 = h ? ((s2 & 0x0F0F0F0Fu) | ((s0 >> 2) & 0x30303030u)) : (s0 & 0x3F3F3F3Fu);
        const uint32_t mn4 = h ? (((s2 >> 4) & 0x0F0F0F0Fu) | ((s1 >> 2) & 0x30303030u)) : (s1 & 0x3F3F3F3Fu);
        float fd = 0.0f, fm = 0.0f;
#pragma unroll
        for (int g = 0; g < 2; g++) {
            const uint32_t qw[8] = { w.q[2 * g].x, w.q[2 * g].y, w.q[2 * g].z, w.q[2 * g].w,
                                     w.q[2 * g + 1].x, w.q[2 * g + 1].y, w.q[2 * g + 1].z, w.q[2 * g + 1].w };
            int sl = 0, sh = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                sl = __dp4a((int) (qw[i] & 0x0F0F0F0Fu), a.q[2 * g][i], sl);
                sh = dp4a_us(qw[i] & 0xF0F0F0F0u, a.q[2 * g + 1][i], sh);      // 16 x the nibble dot
            }
            const int sc0 = (int) ((sc4 >> (16 * g)) & 0xFFu),     m0 = (int) ((mn4 >> (16 * g)) & 0xFFu);
            const int sc1 = (int) ((sc4 >> (16 * g + 8)) & 0xFFu), m1 = (int) ((mn4 >> (16 * g + 8)) & 0xFFu);
            fd += a.d[2 * g] * (float) (sl * sc0) + a.d[2 * g + 1] * (float) (sh * sc1);
            fm += a.sd[2 * g] * (float) m0 + a.sd[2 * g + 1] * (float) m1;
        }
        return d * fd - dmin * fm;
    }
    static __device__ __forceinline__ float load_dot(const uint8_t * row, int idx, const acts128 & a) {
        wts w; load(w, row, idx); return dot(w, idx & 1, a);
    }
};

// ---- Q5_K: block 176 B = d,dmin | scales[12] | qh[32] | qs[128]; bit j of qh[l] is the 5th bit of
// element l of sub-block j.
template <> struct hitem<T_Q5_K> {
    static constexpr bool aligned16 = true;
    static constexpr bool need_ys16 = false;
    static constexpr bool hi_div16 = false;
    static bool ok(int64_t k, int64_t row_stride, uintptr_t base) { return k % 256 == 0 && row_stride % 16 == 0 && base % 16 == 0; }
    struct wts { uint4 hdr; uint4 qh[2]; uint4 q[4]; };
    static __device__ __forceinline__ void load(wts & w, const uint8_t * row, int idx) {
        const uint8_t * bp = row + (idx >> 1) * 176;
        w.hdr = lds128(bp);
        w.qh[0] = lds128(bp + 16); w.qh[1] = lds128(bp + 32);
        const uint8_t * qp = bp + 48 + 64 * (idx & 1);
#pragma unroll
        for (int i = 0; i < 4; i++) w.q[i] = lds128(qp + 16 * i);
    }
    static __device__ __forceinline__ float dot(const wts & w, int h, const acts128 & a) {
        const float d = f16_bits_to_f32(w.hdr.x), dmin = f16_bits_to_f32(w.hdr.x >> 16);
        const uint32_t s0 = w.hdr.y, s1 = w.hdr.z, s2 = w.hdr.w;
        const uint32_t sc4 = h ? ((s2 & 0x0F0F0F0Fu) | ((s0 >> 2) & 0x30303030u)) : (s0 & 0x3F3F3F3Fu);
        const uint32_t mn4 = h ? (((s2 >> 4) & 0x0F0F0F0Fu) | ((s1 >> 2) & 0x30303030u)) : (s1 & 0x3F3F3F3Fu);
        const uint32_t hb[8] = { w.qh[0].x, w.qh[0].y, w.qh[0].z, w.qh[0].w, w.qh[1].x, w.qh[1].y, w.qh[1].z, w.qh[1].w };
        float fd = 0.0f, fm = 0.0f;
#pragma unroll
        for (int g = 0; g < 2; g++) {
            const uint32_t qw[8] = { w.q[2 * g].x, w.q[2 * g].y, w.q[2 * g].z, w.q[2 * g].w,
                                     w.q[2 * g + 1].x, w.q[2 * g + 1].y, w.q[2 * g + 1].z, w.q[2 * g + 1].w };
            int sl = 0, sh = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint32_t b = hb[i] >> (4 * h + 2 * g);           // bit 0: low group, bit 1: high group
                const uint32_t vl = (qw[i] & 0x0F0F0F0Fu) | ((b << 4) & 0x10101010u);
                const uint32_t vh = ((qw[i] >> 4) & 0x0F0F0F0Fu) | ((b << 3) & 0x10101010u);
                sl = __dp4a((int) vl, a.q[2 * g][i], sl);
                sh = __dp4a((int) vh, a.q[2 * g + 1][i], sh);
            }
            const int sc0 = (int) ((sc4 >> (16 * g)) & 0xFFu),     m0 = (int) ((mn4 >> (16 * g)) & 0xFFu);
            const int sc1 = (int) ((sc4 >> (16 * g + 8)) & 0xFFu), m1 = (int) ((mn4 >> (16 * g + 8)) & 0xFFu);
            fd += a.d[2 * g] * (float) (sl * sc0) + a.d[2 * g + 1] * (float) (sh * sc1);
            fm += a.sd[2 * g] * (float) m0 + a.sd[2 * g + 1] * (float) m1;
        }
        return d * fd - dmin * fm;
    }
    static __device__ __forceinline__ float load_dot(const uint8_t * row, int idx, const acts128 & a) {
        wts w; load(w, row, idx); return dot(w, idx & 1, a);
    }
};

// ---- Q6_K: block 210 B = ql[128] | qh[64] | int8 scales[16] | f16 d (src/ggml-common.h:330-336); item h of
// block b = elements 128h .. 128h+127 = four streams of 32 (src/ggml-quants.c:1762-1791):
//   stream 0: (ql[64h+l]    & 15) | ((qh[32h+l]     ) & 3) << 4      stream 2: (ql[64h+l]    >> 4) | ((qh[32h+l] >> 4) & 3) << 4
//   stream 1: (ql[64h+32+l] & 15) | ((qh[32h+l] >> 2) & 3) << 4      stream 3: (ql[64h+32+l] >> 4) | ((qh[32h+l] >> 6) & 3) << 4
// w = d * scales[8h + 2s + l/16] * (q - 32).  Blocks are only 2-byte aligned: every run of words is fetched as
// aligned words + funnel shift.  The -32 is taken out of the dot product through the activation sums per 16.
template <> struct hitem<T_Q6_K> {
    static constexpr bool aligned16 = false;
    static constexpr bool need_ys16 = true;
    static constexpr bool hi_div16 = false;
    static bool ok(int64_t k, int64_t row_stride, uintptr_t base) { return k % 256 == 0 && row_stride % 2 == 0 && base % 2 == 0; }
    struct wts { uint32_t ql[16]; uint32_t qh[8]; uint32_t sc[2]; uint32_t d; };
    template <int N>
    static __device__ __forceinline__ void run(uint32_t (&out)[N], const uint8_t * p) {      // N words from a 2-byte aligned address
        const uintptr_t a = reinterpret_cast<uintptr_t>(p);
        const uint32_t * A = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t) 3);
        const uint32_t sh = (uint32_t) (a & 2) * 8;
        uint32_t x[N + 1];
#pragma unroll
        for (int i = 0; i <= N; i++) x[i] = A[i];
#pragma unroll
        for (int i = 0; i < N; i++) out[i] = __funnelshift_r(x[i], x[i + 1], sh);
    }
    static __device__ __forceinline__ void load(wts & w, const uint8_t * row, int idx) {
        const uint8_t * bp = row + (idx >> 1) * 210;
        const int h = idx & 1;
        run<16>(w.ql, bp + 64 * h);
        run<8>(w.qh, bp + 128 + 32 * h);
        run<2>(w.sc, bp + 192 + 8 * h);
        w.d = lds16(bp + 208);
    }
    static __device__ __forceinline__ float dot(const wts & w, int, const acts128 & a) {
        float r = 0.0f;
#pragma unroll
        for (int s = 0; s < 4; s++) {
            int sum[2] = { 0, 0 };
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint32_t l = w.ql[8 * (s & 1) + i];
                const uint32_t lo = ((s & 2) ? (l >> 4) : l) & 0x0F0F0F0Fu;
                const uint32_t hb = s == 0 ? (w.qh[i] << 4) : (s == 1 ? (w.qh[i] << 2) : (s == 2 ? w.qh[i] : (w.qh[i] >> 2)));
                sum[i >> 2] = __dp4a((int) (lo | (hb & 0x30303030u)), a.q[s][i], sum[i >> 2]);
            }
            // scales 2s, 2s+1 of this half: bytes of sc[s >> 1]
            const uint32_t sw = w.sc[s >> 1] >> (16 * (s & 1));
            const int sc0 = (int) (int8_t) (sw & 0xFFu), sc1 = (int) (int8_t) ((sw >> 8) & 0xFFu);
            r += a.d[s] * (float) (sc0 * (sum[0] - 32 * a.ys16[2 * s]) + sc1 * (sum[1] - 32 * a.ys16[2 * s + 1]));
        }
        return f16_bits_to_f32(w.d) * r;
    }
    // load and reduce in one pass, 8 ql words at a time (streams s and s + 2 share them): half the live registers of
    // load() + dot(), same arithmetic in the same order
    static __device__ __forceinline__ float load_dot(const uint8_t * row, int idx, const acts128 & a) {
        const uint8_t * bp = row + (idx >> 1) * 210;
        const int h = idx & 1;
        uint32_t qh[8], sc[2];
        run<8>(qh, bp + 128 + 32 * h);
        run<2>(sc, bp + 192 + 8 * h);
        float part[4];
#pragma unroll
        for (int half = 0; half < 2; half++) {
            uint32_t ql[8];
            run<8>(ql, bp + 64 * h + 32 * half);
#pragma unroll
            for (int hi = 0; hi < 2; hi++) {
                const int s = half + 2 * hi;
                int sum[2] = { 0, 0 };
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const uint32_t lo = (hi ? (ql[i] >> 4) : ql[i]) & 0x0F0F0F0Fu;
                    const uint32_t hb = s == 0 ? (qh[i] << 4) : (s == 1 ? (qh[i] << 2) : (s == 2 ? qh[i] : (qh[i] >> 2)));
                    sum[i >> 2] = __dp4a((int) (lo | (hb & 0x30303030u)), a.q[s][i], sum[i >> 2]);
                }
                const uint32_t sw = sc[s >> 1] >> (16 * (s & 1));
                const int sc0 = (int) (int8_t) (sw & 0xFFu), sc1 = (int) (int8_t) ((sw >> 8) & 0xFFu);
                part[s] = a.d[s] * (float) (sc0 * (sum[0] - 32 * a.ys16[2 * s]) + sc1 * (sum[1] - 32 * a.ys16[2 * s + 1]));
            }
        }
        const float r = ((part[0] + part[1]) + part[2]) + part[3];      // the order dot() adds them in
        return f16_bits_to_f32(lds16(bp + 208)) * r;
    }
};

// ---- Q4_0 / Q8_0: item = four 32-weight blocks (72 / 136 bytes).  Blocks are 18 / 34 bytes: block b of an item
// starts at word 4.5 b / 8.5 b, so even blocks have their f16 d in the low half of a word and their quants across
// word boundaries, odd blocks the other way round; the item itself is only 2-byte aligned (run<> realigns).
// load() only notes the address; dot() streams block by block so that at most 10 words are live.
template <> struct hitem<T_Q4_0> {
    static constexpr bool aligned16 = false;
    static constexpr bool need_ys16 = false;
    static constexpr bool hi_div16 = false;
    static bool ok(int64_t k, int64_t row_stride, uintptr_t base) { return k % 128 == 0 && row_stride % 2 == 0 && base % 2 == 0; }
    struct wts { const uint8_t * p; };
    static __device__ __forceinline__ void load(wts & w, const uint8_t * row, int idx) { w.p = row + idx * 72; }
    static __device__ __forceinline__ float dot(const wts & w, int, const acts128 & a) {
        // the item's 18 words: 8-byte loads when the item is 8-byte aligned (items are 72 bytes apart: true for every
        // lane as soon as the row start is), else aligned words + funnel shift
        uint32_t x[18];
        if ((reinterpret_cast<uintptr_t>(w.p) & 7) == 0) {
#pragma unroll
            for (int i = 0; i < 9; i++) {
                const uint2 v = *reinterpret_cast<const uint2 *>(w.p + 8 * i);
                x[2 * i] = v.x; x[2 * i + 1] = v.y;
            }
        } else {
            hitem<T_Q6_K>::run<18>(x, w.p);
        }
        float r = 0.0f;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            constexpr int B0[4] = { 0, 4, 9, 13 };
            uint32_t d, q[4];
            if ((b & 1) == 0) {
                d = x[B0[b]] & 0xFFFFu;
#pragma unroll
                for (int i = 0; i < 4; i++) q[i] = __funnelshift_r(x[B0[b] + i], x[B0[b] + i + 1], 16);
            } else {
                d = x[B0[b]] >> 16;
#pragma unroll
                for (int i = 0; i < 4; i++) q[i] = x[B0[b] + 1 + i];
            }
            int lo = 0, hi16 = 0;                       // hi16 = 16 x the high-nibble dot (nibbles used in place)
#pragma unroll
            for (int i = 0; i < 4; i++) {
                lo   = __dp4a((int) (q[i] & 0x0F0F0F0Fu), a.q[b][i], lo);
                hi16 = dp4a_us(q[i] & 0xF0F0F0F0u, a.q[b][4 + i], hi16);
            }
            const int sum = lo + (hi16 >> 4);           // exact: hi16 is a multiple of 16
            r += f16_bits_to_f32(d) * (a.d[b] * (float) sum - 8.0f * a.sd[b]);
        }
        return r;
    }
};

template <> struct hitem<T_Q8_0> {
    static constexpr bool aligned16 = false;
    static constexpr bool need_ys16 = false;
    static constexpr bool hi_div16 = false;
    static bool ok(int64_t k, int64_t row_stride, uintptr_t base) { return k % 128 == 0 && row_stride % 2 == 0 && base % 2 == 0; }
    struct wts { const uint8_t * p; };
    static __device__ __forceinline__ void load(wts & w, const uint8_t * row, int idx) { w.p = row + idx * 136; }
    static __device__ __forceinline__ float dot(const wts & w, int, const acts128 & a) {
        float r = 0.0f;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            constexpr int B0[4] = { 0, 8, 17, 25 };
            uint32_t x[9];
            hitem<T_Q6_K>::run<9>(x, w.p + 4 * B0[b]);
            uint32_t d, q[8];
            if ((b & 1) == 0) {
                d = x[0] & 0xFFFFu;
#pragma unroll
                for (int i = 0; i < 8; i++) q[i] = __funnelshift_r(x[i], x[i + 1], 16);
            } else {
                d = x[0] >> 16;
#pragma unroll
                for (int i = 0; i < 8; i++) q[i] = x[1 + i];
            }
            int sum = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) sum = __dp4a((int) q[i], a.q[b][i], sum);
            r += f16_bits_to_f32(d) * (a.d[b] * (float) sum);
        }
        return r;
    }
};

// ================================================================= the kernel
struct fused_args {
    const uint8_t * w;        // weight rows
    int64_t row_stride;       // bytes between rows
    int64_t m;                // rows
    const float * x;          // activations, k floats, 16-byte aligned
    float * dst;              // m floats
    int     k;
    int     row_bytes;        // bytes of one row
    int     items;            // 128-weight items per row (k / 128)
    int     rows_per_cta;
    int     R, S;             // rows per stage, stages
    int     W, G;             // warps per row group, row groups per CTA (consumer warps = W * G)
    int     slot_stride;      // bytes per stage slot (multiple of 128)
    int     off_part, off_yq, off_yd, off_ys, off_slots;   // shared-memory byte offsets
    int     contiguous;       // rows back to back in memory -> one bulk copy per stage
    int     row_pitch;        // non-contiguous: bytes between rows inside a slot (multiple of 16)
    int     early_weights;    // weights are constant: request them before the grid dependency resolves
#ifdef B200_TUNING
    int     dbg_mode;         // tuning aid: 1 = skip the dot products, 2 = skip the quantization
    unsigned long long * dbg; // tuning aid: per-CTA globaltimer stamps (NULL in normal use)
#endif
};

constexpr int FUSED_MAX_WARPS = 10;   // up to 9 consumer warps + the producer
constexpr int FUSED_MAX_STAGES = 16;

template <int T, int IPL>
__global__ void __launch_bounds__(FUSED_MAX_WARPS * 32)
k_mmvq_fused(const fused_args a) {
    using IT = hitem<T>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t * full  = reinterpret_cast<uint64_t *>(smem);
    uint64_t * empty = full + FUSED_MAX_STAGES;
    float    * part  = reinterpret_cast<float *>(smem + a.off_part);
    int8_t   * yq    = reinterpret_cast<int8_t *>(smem + a.off_yq);
    float    * yd    = reinterpret_cast<float *>(smem + a.off_yd);
    int      * ys    = reinterpret_cast<int *>(smem + a.off_ys);
    uint8_t  * slots = smem + a.off_slots;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int W = a.W, G = a.G, NWC = W * G;
    const int R = a.R, S = a.S;
    const int64_t r_begin = (int64_t) blockIdx.x * a.rows_per_cta;
    const int nrows = (int) ((a.m - r_begin < a.rows_per_cta) ? (a.m - r_begin) : a.rows_per_cta);
    const int nbatch = (nrows + R - 1) / R;

    if (tid == 0) {
        for (int s = 0; s < S; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], (uint32_t) NWC); }
        mbar_fence_init();
    }
    __syncthreads();
    pdl_launch();

    if (warp == NWC) {
        // ===================================================== producer: weight rows -> stage ring
        // The WHOLE warp waits on the empty barrier and lane 0 issues: an mbarrier wait executed by a
        // single lane is only released when its suspend time limit runs out (~0.4 us measured with
        // tools/stream_probe.cu), which throttles the ring to one slot per time-out.
        if (lane == 0) FUSED_STAMP(0);
        if (!a.early_weights) pdl_wait();
        for (int b = 0; b < nbatch; b++) {
            const int s = b % S;
            if (b >= S) mbar_wait(&empty[s], (uint32_t) (((b / S) - 1) & 1));
            if (lane == 0) {
                const int rows_here = (nrows - b * R < R) ? (nrows - b * R) : R;
                const uint8_t * src = a.w + (r_begin + (int64_t) b * R) * a.row_stride;
                uint8_t * slot = slots + (size_t) s * a.slot_stride;
#ifdef B200_TUNING
                if (a.dbg && blockIdx.x == 0 && b < 120) a.dbg[2048 + b] = gtime();
#endif
                if (a.contiguous) {
                    const uint32_t off = (uint32_t) (reinterpret_cast<uintptr_t>(src) & 15);
                    const uint32_t bytes = (off + (uint32_t) rows_here * (uint32_t) a.row_bytes + 15u) & ~15u;
                    mbar_arrive_expect_tx(&full[s], bytes);
                    bulk_g2s(slot, src - off, bytes, &full[s]);
                } else {
                    uint32_t total = 0;
                    for (int r = 0; r < rows_here; r++) {
                        const uint32_t off = (uint32_t) (reinterpret_cast<uintptr_t>(src + r * a.row_stride) & 15);
                        total += (off + (uint32_t) a.row_bytes + 15u) & ~15u;
                    }
                    mbar_arrive_expect_tx(&full[s], total);
                    for (int r = 0; r < rows_here; r++) {
                        const uint8_t * p = src + r * a.row_stride;
                        const uint32_t off = (uint32_t) (reinterpret_cast<uintptr_t>(p) & 15);
                        bulk_g2s(slot + r * a.row_pitch, p - off, (off + (uint32_t) a.row_bytes + 15u) & ~15u, &full[s]);
                    }
                }
            }
            __syncwarp();
        }
        return;
    }

    // ========================================================= consumers
    // 1. quantize x (bit-exact quantize_row_q8_1_ref values) into shared memory: 8 lanes per q8 block
    pdl_wait();
    if (warp == 0) FUSED_STAMP(1);
    if (!FUSED_DBG_MODE(2)) {
        const int nblk32 = a.k >> 5;
        const int sub = lane >> 3, l8 = lane & 7;
        constexpr int U = 4;                                  // loads in flight per lane
        for (int blk0 = warp * 4 + sub; blk0 < nblk32; blk0 += NWC * 4 * U) {
            float4 vv[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int blk = blk0 + u * NWC * 4;
                if (blk < nblk32) vv[u] = __ldg(reinterpret_cast<const float4 *>(a.x + blk * 32 + l8 * 4));
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int blk = blk0 + u * NWC * 4;
                if (blk >= nblk32) break;                     // warp-uniform: k % 128 == 0
                const float4 v = vv[u];
                float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, o));
                const float d  = __fdiv_rn(amax, 127.0f);
                const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
                const float xs[4] = { v.x, v.y, v.z, v.w };
                uint32_t packed = 0;
                int sum = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    // float -> int8 as quantize.cu does it (cvttss2si semantics of the reference's x86 build)
                    const float r  = roundf(__fmul_rn(xs[j], id));
                    const int   qi = fabsf(r) < 2147483648.0f ? (int) r : (int) 0x80000000;
                    const int   q  = (int) (int8_t) (qi & 0xFF);
                    packed |= (uint32_t) (q & 0xFF) << (8 * j);
                    sum += q;
                }
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
                *reinterpret_cast<uint32_t *>(yq + blk * 32 + l8 * 4) = packed;
                if (l8 == 0) { yd[blk] = __half2float(__float2half_rn(d)); ys[blk] = sum; }
            }
        }
    }
    named_bar_sync(1, NWC * 32);
    if (warp == 0) FUSED_STAMP(2);

    // 2. the slice of K this lane owns, in registers
    const int g = warp / W, wg = warp - g * W;
    acts128 acts[IPL];
#pragma unroll
    for (int i = 0; i < IPL; i++) {
        const int idx = (i * W + wg) * 32 + lane;
        if (idx < a.items) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int4 lo = *reinterpret_cast<const int4 *>(yq + idx * 128 + 32 * j);
                const int4 hi = *reinterpret_cast<const int4 *>(yq + idx * 128 + 32 * j + 16);
                acts[i].q[j][0] = lo.x; acts[i].q[j][1] = lo.y; acts[i].q[j][2] = lo.z; acts[i].q[j][3] = lo.w;
                acts[i].q[j][4] = hi.x; acts[i].q[j][5] = hi.y; acts[i].q[j][6] = hi.z; acts[i].q[j][7] = hi.w;
                const float d = yd[idx * 4 + j];
                acts[i].sd[j] = d * (float) ys[idx * 4 + j];
                acts[i].d[j]  = (IT::hi_div16 && (j & 1)) ? d * 0.0625f : d;
            }
        }
    }

    // 3. stream the rows
    int work_it = 0;                  // stages in which this group had rows: parity selects the partials buffer
    for (int b = 0; b < nbatch; b++) {
        const int s = b % S;
        mbar_wait(&full[s], (uint32_t) ((b / S) & 1));      // every lane waits (see the producer's note)
        if (warp == 0 && b == 0) FUSED_STAMP(3);
#ifdef B200_TUNING
        if (a.dbg && blockIdx.x == 0 && warp == 0 && lane == 0 && b < 120) a.dbg[2048 + 128 + b] = gtime();
#endif
        if (warp == 0 && b == nbatch - 1) FUSED_STAMP(4);
        const int rows_here = (nrows - b * R < R) ? (nrows - b * R) : R;
        const uint8_t * slot = slots + (size_t) s * a.slot_stride;
        const uint8_t * src0 = a.w + (r_begin + (int64_t) b * R) * a.row_stride;
        const int nchunks = (rows_here + FUSED_RB - 1) / FUSED_RB;
        float * pb = part + (size_t) (g * 2 + (work_it & 1)) * R * W;      // per group: groups work on different slots at once
        // chunks (row pairs) are dealt to the groups round-robin over the CTA's whole slab, not per slot
        const int c_first = (g - (b * (R / FUSED_RB)) % G + G) % G;

        for (int c = c_first; c < nchunks && !FUSED_DBG_MODE(1); c += G) {
            const int r0 = c * FUSED_RB;
            float acc[FUSED_RB];
#pragma unroll
            for (int rr = 0; rr < FUSED_RB; rr++) acc[rr] = 0.0f;
#pragma unroll
            for (int i = 0; i < IPL; i++) {
                const int idx = (i * W + wg) * 32 + lane;
                if (idx < a.items) {
                    typename IT::wts w[FUSED_RB];
#pragma unroll
                    for (int rr = 0; rr < FUSED_RB; rr++) {
                        const int r = (r0 + rr < rows_here) ? (r0 + rr) : (rows_here - 1);   // odd tail: recompute, never stored
                        const uint8_t * row = a.contiguous
                            ? slot + (reinterpret_cast<uintptr_t>(src0) & 15) + (size_t) r * a.row_bytes
                            : slot + (size_t) r * a.row_pitch + (reinterpret_cast<uintptr_t>(src0 + r * a.row_stride) & 15);
                        IT::load(w[rr], row, idx);
                    }
#pragma unroll
                    for (int rr = 0; rr < FUSED_RB; rr++) acc[rr] += IT::dot(w[rr], idx & 1, acts[i]);
                }
            }
            // one shuffle tree for the row pair: lanes 0-15 end with row r0, lanes 16-31 with row r0 + 1
            float v = (lane & 16) ? acc[1] : acc[0];
            const float o = (lane & 16) ? acc[0] : acc[1];
            v += __shfl_xor_sync(0xFFFFFFFFu, o, 16);
#pragma unroll
            for (int off = 8; off > 0; off >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, off);
            const int r = r0 + (lane >> 4);
            if ((lane & 15) == 0 && r < rows_here) {
                if (W == 1) a.dst[r_begin + (int64_t) b * R + r] = v;
                else        pb[r * W + wg] = v;
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_cta(&empty[s]);          // this warp is done with the slot
#ifdef B200_TUNING
        if (a.dbg && blockIdx.x == 0 && warp == NWC - 1 && lane == 0 && b < 120) a.dbg[2048 + 256 + b] = gtime();
#endif
        if (c_first < nchunks) work_it++;
        if (W > 1 && c_first < nchunks) {
            // the W warps of the group meet; warp 0 of the group adds the partials in fixed order.  The
            // buffer written two working stages later is only touched after the barrier in between,
            // which warp 0 reaches after these reads.
            named_bar_sync(2 + g, W * 32);
            if (wg == 0) {
                for (int j = lane; ; j += 32) {
                    const int c = c_first + (j / FUSED_RB) * G;
                    const int r = c * FUSED_RB + (j % FUSED_RB);
                    if (c >= nchunks) break;
                    if (r < rows_here) {
                        float v = 0.0f;
                        for (int u = 0; u < W; u++) v += pb[r * W + u];
                        a.dst[r_begin + (int64_t) b * R + r] = v;
                    }
                }
            }
        }
    }
    if (warp == 0) FUSED_STAMP(5);
}

} // namespace b200

namespace b200 {

// ================================================================= host side
#ifdef B200_TUNING
static unsigned long long * g_fused_dbg = nullptr;   // set through b200_debug_timestamps (tuning builds only)
#endif

template <int T, int IPL>
static int launch_fused(fused_args a, cudaStream_t st) {
    const dev_caps & caps = current_caps();
    static const int ring_env = env_int("GGML_B200_FUSED_RING_BYTES", 98304);   // two kernels' CTAs share an SM
    static const int slot_env = env_int("GGML_B200_FUSED_SLOT_BYTES", 9216);
    static const int cps_env  = env_int("GGML_B200_FUSED_CTAS_PER_SM", 1);
    static const int pdl      = env_int("GGML_B200_FUSED_PDL", 1);

    int mult = slot_env / (FUSED_RB * a.row_bytes);
    if (mult < 1) mult = 1;
    a.R = FUSED_RB * mult;
    a.row_pitch = (a.row_bytes + 15 + 15 + 16) & ~15;
    const int slot_bytes = a.contiguous ? (15 + a.R * a.row_bytes + 15 + 16) : a.R * a.row_pitch;
    a.slot_stride = (slot_bytes + 127) & ~127;
    a.off_part  = 2 * FUSED_MAX_STAGES * 8;
    a.off_yq    = (a.off_part + a.G * 2 * a.R * a.W * (int) sizeof(float) + 15) & ~15;
    a.off_yd    = a.off_yq + a.k;
    a.off_ys    = a.off_yd + (a.k / 32) * 4;
    a.off_slots = (a.off_ys + (a.k / 32) * 4 + 127) & ~127;
    a.S = ring_env / a.slot_stride;
    if (a.S < 2) a.S = 2;
    if (a.S > FUSED_MAX_STAGES) a.S = FUSED_MAX_STAGES;
    const int cps = cps_env < 1 ? 1 : cps_env;
    const size_t smem = (size_t) a.off_slots + (size_t) a.S * a.slot_stride;
    if (smem > (size_t) caps.smem_optin) return B200_STREAM_NOT_APPLICABLE;

    auto kern = k_mmvq_fused<T, IPL>;
    static smem_limit lim;
    if (const int rc = lim.ensure(kern, smem)) return rc;
    const int dev = current_device_slot();
    int64_t grid = (int64_t) caps.sm_count * cps;
    const int64_t max_grid = ceil_div(a.m, FUSED_RB);
    if (grid > max_grid) grid = max_grid;
    a.rows_per_cta = (int) ceil_div(a.m, grid);
    grid = ceil_div(a.m, a.rows_per_cta);

    static std::atomic<bool> carve_set[64];
    if (!carve_set[dev].load(std::memory_order_acquire)) {
        // two kernels' CTAs must fit an SM at once (the next mat-vec prefetches while this one computes)
        cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        carve_set[dev].store(true, std::memory_order_release);
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) grid); cfg.blockDim = dim3((unsigned) ((a.W * a.G + 1) * 32));
    cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, a);
    if (e != cudaSuccess) return -(int) e;
    b200_count_launch(1);
    return B200_OK;
}

template <int T>
static int launch_fused_t(fused_args a, cudaStream_t st) {
    if (!hitem<T>::ok(a.k, a.row_stride, reinterpret_cast<uintptr_t>(a.w))) return B200_STREAM_NOT_APPLICABLE;
    a.items = a.k / 128;
    const int passes = (int) ceil_div(a.items, 32);             // warp-passes one row needs
    if (passes <= 8) {
        a.W = passes; a.G = 9 / passes;                         // up to 9 consumer warps
        return launch_fused<T, 1>(a, st);
    }
    if (passes <= 16) {
        a.W = (passes + 1) / 2; a.G = 1;
        return launch_fused<T, 2>(a, st);
    }
    return B200_STREAM_NOT_APPLICABLE;
}

// dst[m] = W[m, k] . x[k]; returns B200_OK, B200_STREAM_NOT_APPLICABLE (caller falls back) or an error.
static int mmvq_fused_try(int type, const uint8_t * w, int64_t row_stride, int64_t m, int64_t k,
                          const float * x, float * dst, bool const_weights, cudaStream_t st) {
    if (b200_option(B200_OPT_MMVQ_NO_FUSED)) return B200_STREAM_NOT_APPLICABLE;
    if (m < 1 || k < 1024 || k % 256 != 0 || k > 131072) return B200_STREAM_NOT_APPLICABLE;
    if (reinterpret_cast<uintptr_t>(x) % 16 != 0) return B200_STREAM_NOT_APPLICABLE;
    const int64_t row_bytes = k / type_qk(type) * type_bytes(type);
    if (row_stride < row_bytes) return B200_STREAM_NOT_APPLICABLE;
    fused_args a = {};
    a.w = w; a.row_stride = row_stride; a.m = m; a.x = x; a.dst = dst; a.k = (int) k; a.row_bytes = (int) row_bytes;
    a.contiguous = row_stride == row_bytes ? 1 : 0;
    a.early_weights = const_weights ? 1 : 0;
#ifdef B200_TUNING
    a.dbg = g_fused_dbg;
    a.dbg_mode = env_int("GGML_B200_FUSED_DBG_MODE", 0);
#endif
    switch (type) {
        case T_Q4_K: return launch_fused_t<T_Q4_K>(a, st);
        case T_Q5_K: return launch_fused_t<T_Q5_K>(a, st);
        default: return B200_STREAM_NOT_APPLICABLE;
    }
}

} // namespace b200
