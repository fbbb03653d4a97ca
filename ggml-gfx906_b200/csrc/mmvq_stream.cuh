// mmvq_stream.cuh — the B200 decode mat-vec: TMA-staged, K-stationary weight streaming.
//
// Design (what is different from the reference's mul_mat_vec_q, src/ggml-cuda/mmvq.cu:139-226):
//   * The CTA covers the WHOLE reduction dimension: every lane owns a fixed slice of K ("item") and
//     keeps the matching q8 activations (and their sums / scales) in registers for the lifetime of
//     the kernel.  Activations are read once per CTA, never again.
//   * Weight rows are streamed through a ring of shared-memory stages by bulk async copies
//     (cp.async.bulk = TMA 1-D, SASS UBLKCP) issued by one thread and completed on mbarriers, so the
//     number of bytes in flight per SM is set by the ring depth (tens of KiB), not by occupancy or
//     per-thread loads.  Copies are 16-byte-aligned supersets of the rows, which removes the
//     2-byte-alignment problem of the 18/34/210-byte block formats from the global-memory side.
//   * Consumers read the staged rows with bank-conflict-free word patterns chosen per format
//     (Q4_K/Q5_K: one 16-byte quant chunk per lane; Q4_0/Q8_0: one block PAIR per lane, 9/17-word
//     stride is odd; Q6_K: 16 lanes per block).
//   * Per row: integer dp4a dots -> f32, warp-shuffle reduction, per-warp partial to shared memory,
//     one named barrier per stage, a rotating warp adds the partials in fixed order (deterministic:
//     the result of a row never depends on the grid or on which GPU computes it).
//
// Grid: one CTA per contiguous slab of rows, sized so that the grid is a multiple of the SM count.
#pragma once
#include <atomic>
#include <cstdlib>
#include "common.cuh"

namespace b200 {

constexpr int B200_STREAM_NOT_APPLICABLE = B200_NOT_APPLICABLE;

// ---------------------------------------------------------------- PTX helpers (mbarrier + bulk copy)
__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t * bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t * bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t * bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "B200_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra B200_DONE;\n"
        "bra B200_WAIT;\n"
        "B200_DONE:\n"
        "}\n" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy; dst, src 16-byte aligned, bytes % 16 == 0
__device__ __forceinline__ void bulk_g2s(void * dst, const void * src, uint32_t bytes, uint64_t * bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(nthreads) : "memory");
}

// aligned shared-memory word loads
__device__ __forceinline__ uint32_t lds32(const uint8_t * p) { return *reinterpret_cast<const uint32_t *>(p); }
__device__ __forceinline__ uint4    lds128(const uint8_t * p) { return *reinterpret_cast<const uint4 *>(p); }
// 32 bits from a 2-byte aligned shared address: two aligned words + funnel shift (branch-free)
__device__ __forceinline__ uint32_t lds32_a2(const uint8_t * p) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t * w = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t) 3);
    return __funnelshift_r(w[0], w[1], (uint32_t) (a & 2) * 8);
}
__device__ __forceinline__ uint32_t lds16(const uint8_t * p) { return (uint32_t) *reinterpret_cast<const uint16_t *>(p); }

// ================================================================= per-format items
// An "item" is the slice of one weight row a lane handles per row.  Each format defines
//   WPI            weights per item
//   struct acts    the register-resident activation data of one item for one column
//   load_acts()    fill it from the q8 planes (once per kernel)
//   dot()          partial dot product of the item of the staged row with acts
//   ok()           whether the row geometry suits the item mapping
template <int T> struct item;

// ---- Q4_K: item idx -> block idx/8, 16-byte quant chunk j = idx%8 (64-elem group g=j/2, half h=j%2):
// low nibbles = elements 64g+16h+[0,16) (sub-block 2g), high nibbles = +32 (sub-block 2g+1).
template <> struct item<T_Q4_K> {
    static constexpr int WPI = 32;
    struct acts { int yl[4], yh[4]; float dl, dh, syl, syh; };
    static bool ok(int64_t k, int64_t row_stride, uintptr_t base) { return k % 256 == 0 && row_stride % 16 == 0 && base % 16 == 0; }
    static __device__ __forceinline__ void load_acts(acts & a, const int8_t * yq, const float * yd, int idx) {
        const int blk = idx >> 3, j = idx & 7, g = j >> 1, h = j & 1;
        const int4 l = *reinterpret_cast<const int4 *>(yq + blk * 256 + 64 * g + 16 * h);
        const int4 u = *reinterpret_cast<const int4 *>(yq + blk * 256 + 64 * g + 16 * h + 32);
        a.yl[0] = l.x; a.yl[1] = l.y; a.yl[2] = l.z; a.yl[3] = l.w;
        a.yh[0] = u.x; a.yh[1] = u.y; a.yh[2] = u.z; a.yh[3] = u.w;
        a.dl = yd[blk * 8 + 2 * g]; a.dh = yd[blk * 8 + 2 * g + 1];
        int sl = 0, sh = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) { sl = __dp4a(0x01010101, a.yl[i], sl); sh = __dp4a(0x01010101, a.yh[i], sh); }
        a.syl = a.dl * (float) sl; a.syh = a.dh * (float) sh;
    }
    struct wts { uint4 hdr, q; int j; };
    static __device__ __forceinline__ void load_w(wts & w, const uint8_t * row, int idx) {
        const uint8_t * bp = row + (idx >> 3) * 144;
        w.j = idx & 7;
        w.hdr = lds128(bp);
        w.q   = lds128(bp + 16 + 16 * w.j);
    }
    static __device__ __forceinline__ void scales(const wts & w, float & d, float & dmin, int & sc0, int & m0, int & sc1, int & m1) {
        d    = f16_bits_to_f32(w.hdr.x);
        dmin = f16_bits_to_f32(w.hdr.x >> 16);
        const int g = w.j >> 1;
        k4_scale_min(w.hdr.y, w.hdr.z, w.hdr.w, 2 * g,     sc0, m0);
        k4_scale_min(w.hdr.y, w.hdr.z, w.hdr.w, 2 * g + 1, sc1, m1);
    }
    static __device__ __forceinline__ float dot(const wts & w, const acts & a) {
        float d, dmin; int sc0, m0, sc1, m1;
        scales(w, d, dmin, sc0, m0, sc1, m1);
        const uint32_t q[4] = { w.q.x, w.q.y, w.q.z, w.q.w };
        int sl = 0, sh = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            sl = __dp4a((int) (q[i] & 0x0F0F0F0Fu), a.yl[i], sl);
            sh = __dp4a((int) ((q[i] >> 4) & 0x0F0F0F0Fu), a.yh[i], sh);
        }
        return d * (a.dl * (float) (sl * sc0) + a.dh * (float) (sh * sc1)) - dmin * (a.syl * (float) m0 + a.syh * (float) m1);
    }
};

// ---- Q5_K: as Q4_K plus the 5th bits: qh[16h + i] bit 2g (low group) / 2g+1 (high group).
template <> struct item<T_Q5_K> {
    static constexpr int WPI = 32;
    using acts = item<T_Q4_K>::acts;
    static bool ok(int64_t k, int64_t row_stride, uintptr_t base) { return k % 256 == 0 && row_stride % 16 == 0 && base % 16 == 0; }
    static __device__ __forceinline__ void load_acts(acts & a, const int8_t * yq, const float * yd, int idx) {
        item<T_Q4_K>::load_acts(a, yq, yd, idx);
    }
    struct wts { uint4 hdr, q, qh; int j; };
    static __device__ __forceinline__ void load_w(wts & w, const uint8_t * row, int idx) {
        const uint8_t * bp = row + (idx >> 3) * 176;
        w.j = idx & 7;
        w.hdr = lds128(bp);
        w.qh  = lds128(bp + 16 + 16 * (w.j & 1));
        w.q   = lds128(bp + 48 + 16 * w.j);
    }
    static __device__ __forceinline__ float dot(const wts & w, const acts & a) {
        const float d    = f16_bits_to_f32(w.hdr.x);
        const float dmin = f16_bits_to_f32(w.hdr.x >> 16);
        const int g = w.j >> 1;
        int sc0, m0, sc1, m1;
        k4_scale_min(w.hdr.y, w.hdr.z, w.hdr.w, 2 * g,     sc0, m0);
        k4_scale_min(w.hdr.y, w.hdr.z, w.hdr.w, 2 * g + 1, sc1, m1);
        const uint32_t q[4]  = { w.q.x, w.q.y, w.q.z, w.q.w };
        const uint32_t qh[4] = { w.qh.x, w.qh.y, w.qh.z, w.qh.w };
        int sl = 0, sh = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t hb = qh[i] >> (2 * g);
            const uint32_t vl = (q[i] & 0x0F0F0F0Fu) | ((hb & 0x01010101u) << 4);
            const uint32_t vh = ((q[i] >> 4) & 0x0F0F0F0Fu) | (((hb >> 1) & 0x01010101u) << 4);
            sl = __dp4a((int) vl, a.yl[i], sl);
            sh = __dp4a((int) vh, a.yh[i], sh);
        }
        return d * (a.dl * (float) (sl * sc0) + a.dh * (float) (sh * sc1)) - dmin * (a.syl * (float) m0 + a.syh * (float) m1);
    }
};

// ---- Q6_K: item idx -> block idx/16, half h = (idx%16)/8, t = idx%8: elements 128h + 32s + 4t + [0,4)
// for the four streams s (16 weights).  Blocks are 210 B: odd blocks sit at 2 mod 4, so every word
// is fetched as two aligned words + funnel shift.
template <> struct item<T_Q6_K> {
    static constexpr int WPI = 16;
    struct acts { int y[4]; float d8[4]; };
    static bool ok(int64_t k, int64_t row_stride, uintptr_t base) { return k % 256 == 0 && row_stride % 2 == 0 && base % 2 == 0; }
    static __device__ __forceinline__ void load_acts(acts & a, const int8_t * yq, const float * yd, int idx) {
        const int blk = idx >> 4, h = (idx >> 3) & 1, t = idx & 7;
#pragma unroll
        for (int s = 0; s < 4; s++) {
            a.y[s]  = *reinterpret_cast<const int *>(yq + blk * 256 + 128 * h + 32 * s + 4 * t);
            a.d8[s] = yd[blk * 8 + 4 * h + s];
        }
    }
    struct wts { uint32_t qla, qlb, qh, s0, s1, d; int t; };
    static __device__ __forceinline__ void load_w(wts & w, const uint8_t * row, int idx) {
        const uint8_t * bp = row + (idx >> 4) * 210;
        const int h = (idx >> 3) & 1;
        w.t = idx & 7;
        w.qla = lds32_a2(bp + 64 * h + 4 * w.t);
        w.qlb = lds32_a2(bp + 64 * h + 32 + 4 * w.t);
        w.qh  = lds32_a2(bp + 128 + 32 * h + 4 * w.t);
        w.s0  = lds32_a2(bp + 192 + 8 * h);
        w.s1  = lds32_a2(bp + 196 + 8 * h);
        w.d   = lds16(bp + 208);
    }
    static __device__ __forceinline__ float dot(const wts & w, const acts & a) {
        const float d = f16_bits_to_f32(w.d);
        const int hi16 = w.t >> 2;                 // l/16 for l = 4t..4t+3
        float r = 0.0f;
#pragma unroll
        for (int s = 0; s < 4; s++) {
            const uint32_t lo = (((s & 1) ? w.qlb : w.qla) >> (4 * (s >> 1))) & 0x0F0F0F0Fu;
            const uint32_t hi = ((w.qh >> (2 * s)) & 0x03030303u) << 4;
            const int v = __vsubss4((int) (lo | hi), 0x20202020);
            const int bi = 2 * s + hi16;           // scale byte 0..7 inside this half
            const int sc = (int) (int8_t) ((((bi < 4) ? w.s0 : w.s1) >> (8 * (bi & 3))) & 0xFFu);
            r += a.d8[s] * (float) (__dp4a(v, a.y[s], 0) * sc);
        }
        return d * r;
    }
};

// ---- Q4_0: item = block pair (36 B = 9 words; lane stride 9 words is conflict-free).  All pairs of a
// row share the same 4-byte phase (36 % 4 == 0), so the phase branch is warp-uniform.
template <> struct item<T_Q4_0> {
    static constexpr int WPI = 64;
    struct acts { int y[2][8]; float d8[2]; int sy8[2]; };
    static bool ok(int64_t k, int64_t row_stride, uintptr_t base) { return k % 64 == 0 && row_stride % 2 == 0 && base % 2 == 0; }
    static __device__ __forceinline__ void load_acts(acts & a, const int8_t * yq, const float * yd, int idx) {
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const int4 l = *reinterpret_cast<const int4 *>(yq + (2 * idx + b) * 32);
            const int4 u = *reinterpret_cast<const int4 *>(yq + (2 * idx + b) * 32 + 16);
            a.y[b][0] = l.x; a.y[b][1] = l.y; a.y[b][2] = l.z; a.y[b][3] = l.w;
            a.y[b][4] = u.x; a.y[b][5] = u.y; a.y[b][6] = u.z; a.y[b][7] = u.w;
            a.d8[b] = yd[2 * idx + b];
            int s = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) s = __dp4a(0x01010101, a.y[b][i], s);
            a.sy8[b] = 8 * s;
        }
    }
    struct wts { uint32_t qs[2][4]; uint32_t d[2]; };
    static __device__ __forceinline__ void load_w(wts & w, const uint8_t * row, int idx) {
        const uint8_t * p = row + idx * 36;
        const uintptr_t a = reinterpret_cast<uintptr_t>(p);
        const uint32_t * A = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t) 3);
        if ((a & 2) == 0) {
            uint32_t x[9];
#pragma unroll
            for (int i = 0; i < 9; i++) x[i] = A[i];
            w.d[0] = x[0] & 0xFFFFu;
#pragma unroll
            for (int i = 0; i < 4; i++) w.qs[0][i] = __funnelshift_r(x[i], x[i + 1], 16);
            w.d[1] = x[4] >> 16;
#pragma unroll
            for (int i = 0; i < 4; i++) w.qs[1][i] = x[5 + i];
        } else {
            uint32_t x[10];
#pragma unroll
            for (int i = 0; i < 10; i++) x[i] = A[i];
            w.d[0] = x[0] >> 16;
#pragma unroll
            for (int i = 0; i < 4; i++) w.qs[0][i] = x[1 + i];
            w.d[1] = x[5] & 0xFFFFu;
#pragma unroll
            for (int i = 0; i < 4; i++) w.qs[1][i] = __funnelshift_r(x[5 + i], x[6 + i], 16);
        }
    }
    static __device__ __forceinline__ float dot(const wts & w, const acts & a) {
        float r = 0.0f;
#pragma unroll
        for (int b = 0; b < 2; b++) {
            int s = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                s = __dp4a((int) (w.qs[b][i] & 0x0F0F0F0Fu), a.y[b][i], s);
                s = __dp4a((int) ((w.qs[b][i] >> 4) & 0x0F0F0F0Fu), a.y[b][i + 4], s);
            }
            r += f16_bits_to_f32(w.d[b]) * a.d8[b] * (float) (s - a.sy8[b]);
        }
        return r;
    }
};

// ---- Q8_0: item = block pair (68 B = 17 words, odd stride); phase handling as Q4_0.
template <> struct item<T_Q8_0> {
    static constexpr int WPI = 64;
    struct acts { int y[2][8]; float d8[2]; };
    static bool ok(int64_t k, int64_t row_stride, uintptr_t base) { return k % 64 == 0 && row_stride % 2 == 0 && base % 2 == 0; }
    static __device__ __forceinline__ void load_acts(acts & a, const int8_t * yq, const float * yd, int idx) {
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const int4 l = *reinterpret_cast<const int4 *>(yq + (2 * idx + b) * 32);
            const int4 u = *reinterpret_cast<const int4 *>(yq + (2 * idx + b) * 32 + 16);
            a.y[b][0] = l.x; a.y[b][1] = l.y; a.y[b][2] = l.z; a.y[b][3] = l.w;
            a.y[b][4] = u.x; a.y[b][5] = u.y; a.y[b][6] = u.z; a.y[b][7] = u.w;
            a.d8[b] = yd[2 * idx + b];
        }
    }
    struct wts { uint32_t qs[2][8]; uint32_t d[2]; };
    static __device__ __forceinline__ void load_w(wts & w, const uint8_t * row, int idx) {
        const uint8_t * p = row + idx * 68;
        const uintptr_t a = reinterpret_cast<uintptr_t>(p);
        const uint32_t * A = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t) 3);
        if ((a & 2) == 0) {
            uint32_t x[17];
#pragma unroll
            for (int i = 0; i < 17; i++) x[i] = A[i];
            w.d[0] = x[0] & 0xFFFFu;
#pragma unroll
            for (int i = 0; i < 8; i++) w.qs[0][i] = __funnelshift_r(x[i], x[i + 1], 16);
            w.d[1] = x[8] >> 16;
#pragma unroll
            for (int i = 0; i < 8; i++) w.qs[1][i] = x[9 + i];
        } else {
            uint32_t x[18];
#pragma unroll
            for (int i = 0; i < 18; i++) x[i] = A[i];
            w.d[0] = x[0] >> 16;
#pragma unroll
            for (int i = 0; i < 8; i++) w.qs[0][i] = x[1 + i];
            w.d[1] = x[9] & 0xFFFFu;
#pragma unroll
            for (int i = 0; i < 8; i++) w.qs[1][i] = __funnelshift_r(x[9 + i], x[10 + i], 16);
        }
    }
    static __device__ __forceinline__ float dot(const wts & w, const acts & a) {
        float r = 0.0f;
#pragma unroll
        for (int b = 0; b < 2; b++) {
            int s = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) s = __dp4a((int) w.qs[b][i], a.y[b][i], s);
            r += f16_bits_to_f32(w.d[b]) * a.d8[b] * (float) s;
        }
        return r;
    }
};

// ================================================================= the kernel
struct stream_args {
    const uint8_t * w;        // weight rows
    int64_t row_stride;       // bytes between rows
    int64_t m;                // rows
    int     row_bytes;        // bytes of one row
    int     items;            // items per row (k / WPI)
    const int8_t * yq;        // activation int8 plane [n][k]
    const float  * yd;        // activation scale plane [n][k/32]
    int64_t k;
    int     n;                // valid columns (<= NCOLS)
    float * dst;
    int64_t dst_col_stride;   // floats
    int     rows_per_cta;
    int     R;                // rows per stage
    int     S;                // stages
    int     slot_stride;      // bytes per stage slot (multiple of 128)
    int     slots_off;        // byte offset of slot 0 in dynamic smem
    int     contiguous;       // rows back to back in memory -> one bulk copy per stage
    int     row_pitch;        // non-contiguous: bytes between rows inside a slot (multiple of 16)
    // MUL_MAT_ID (grid y = slot e, grid z = token t): the weights are expert ids[e, t] of the stack (w_nb2 bytes apart), the
    // activations column t * ne11 + e % ne11 of the planes, the results go to dst + e * dst_nb2 + t * dst_nb3; NULL: plain MUL_MAT
    const uint8_t * ids;
    int64_t ids_nb0, ids_nb1, w_nb2, dst_nb2, dst_nb3;
    int     n_as, ne11;
    unsigned * bad;           // sticky error word: an expert id outside [0, n_as)
    int     n_tok_grid;       // MUL_MAT_ID: tokens (grid z); slots (grid y) = dst_nb3 / dst_nb2
};

template <int NCOLS, int IPL> struct stream_cfg {
    static constexpr int load = NCOLS * IPL;
    static constexpr int max_threads = load <= 2 ? 1024 : (load <= 4 ? 512 : (load <= 8 ? 384 : 192));
};

template <int T, int NCOLS, int IPL>
__global__ void __launch_bounds__(stream_cfg<NCOLS, IPL>::max_threads)
k_mmvq_stream(const stream_args a) {
    using IT = item<T>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t * full = reinterpret_cast<uint64_t *>(smem);
    float    * part = reinterpret_cast<float *>(smem + 64);
    uint8_t  * slots = smem + a.slots_off;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int NW = blockDim.x >> 5;
    const int R = a.R, S = a.S;
    const int64_t r_begin = (int64_t) blockIdx.x * a.rows_per_cta;
    const int nrows = (int) ((a.m - r_begin < a.rows_per_cta) ? (a.m - r_begin) : a.rows_per_cta);
    const int nbatch = (nrows + R - 1) / R;
    const uint8_t * wbase = a.w;
    const int8_t * yq = a.yq;
    const float  * yd = a.yd;
    float * dst = a.dst;
    if (a.ids) {                                              // the same for the whole CTA
        const int64_t i12 = blockIdx.y, i13 = blockIdx.z;
        const int32_t e = *reinterpret_cast<const int32_t *>(a.ids + i12 * a.ids_nb0 + i13 * a.ids_nb1);
        dst += i12 * a.dst_nb2 + i13 * a.dst_nb3;
        if (e < 0 || e >= a.n_as) {                           // the reference asserts (src/ggml-cpu/ggml-cpu.c:1556): flag it, write zeros
            if (tid == 0 && blockIdx.x == 0 && a.bad) atomicOr_system(a.bad, 1u);
            for (int r = tid; r < nrows; r += blockDim.x) dst[r_begin + r] = 0.0f;
            return;
        }
        wbase += (int64_t) e * a.w_nb2;
        const int64_t col0 = i13 * a.ne11 + i12 % a.ne11;
        yq += col0 * a.k; yd += col0 * (a.k / 32);
    }

    auto issue = [&](int b) {
        const int s = b % S;
        const int rows_here = (nrows - b * R < R) ? (nrows - b * R) : R;
        const uint8_t * src = wbase + (r_begin + (int64_t) b * R) * a.row_stride;
        uint8_t * slot = slots + (size_t) s * a.slot_stride;
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
    };

    if (tid == 0) {
        for (int s = 0; s < S; s++) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {
        const int pre = nbatch < S ? nbatch : S;
        for (int b = 0; b < pre; b++) issue(b);
    }

    // register-resident activations (loaded while the first stages are in flight)
    typename IT::acts acts[NCOLS][IPL];
#pragma unroll
    for (int i = 0; i < IPL; i++) {
        const int idx = (i * NW + warp) * 32 + lane;
        if (idx < a.items) {
#pragma unroll
            for (int c = 0; c < NCOLS; c++) {
                const int cc = c < a.n ? c : 0;     // padded columns recompute column 0 (never stored)
                IT::load_acts(acts[c][i], yq + (int64_t) cc * a.k, yd + (int64_t) cc * (a.k / 32), idx);
            }
        }
    }

    for (int b = 0; b < nbatch; b++) {
        const int s = b % S;
        mbar_wait(&full[s], (uint32_t) ((b / S) & 1));
        const int rows_here = (nrows - b * R < R) ? (nrows - b * R) : R;
        const uint8_t * slot = slots + (size_t) s * a.slot_stride;
        const uint8_t * src0 = wbase + (r_begin + (int64_t) b * R) * a.row_stride;
        float * mypart = part + ((size_t) (b & 1) * NW + warp) * R * NCOLS;

        for (int r = 0; r < rows_here; r++) {
            const uint8_t * row = a.contiguous
                ? slot + (reinterpret_cast<uintptr_t>(src0) & 15) + (size_t) r * a.row_bytes
                : slot + (size_t) r * a.row_pitch + (reinterpret_cast<uintptr_t>(src0 + r * a.row_stride) & 15);
            float acc[NCOLS];
#pragma unroll
            for (int c = 0; c < NCOLS; c++) acc[c] = 0.0f;
#pragma unroll
            for (int i = 0; i < IPL; i++) {
                const int idx = (i * NW + warp) * 32 + lane;
                if (idx < a.items) {
                    typename IT::wts w;
                    IT::load_w(w, row, idx);
#pragma unroll
                    for (int c = 0; c < NCOLS; c++) acc[c] += IT::dot(w, acts[c][i]);
                }
            }
#pragma unroll
            for (int c = 0; c < NCOLS; c++) {
                const float v = warp_sum(acc[c]);
                if (lane == 0) mypart[r * NCOLS + c] = v;
            }
        }
        __syncthreads();
        // every warp is past its reads of slot s: refill it
        if (tid == 0 && b + S < nbatch) issue(b + S);
        // a rotating warp adds the per-warp partials in fixed order and writes the rows out
        if (warp == b % NW) {
            const float * pb = part + (size_t) (b & 1) * NW * R * NCOLS;
            for (int o = lane; o < rows_here * NCOLS; o += 32) {
                const int r = o / NCOLS, c = o - r * NCOLS;
                float v = 0.0f;
                for (int w = 0; w < NW; w++) v += pb[(size_t) w * R * NCOLS + o];
                if (c < a.n) dst[(int64_t) c * a.dst_col_stride + r_begin + (int64_t) b * R + r] = v;
            }
        }
    }
}

// ================================================================= host side
// Per-device facts and per-(kernel, device) attribute state are shared by every thread that drives a backend
// (different backends may run on different threads): atomics, no locks on the launch path.
struct dev_caps { int sm_count; int smem_optin; };
inline int current_device_slot() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev < 0 || dev >= 64) ? 0 : dev;
}
inline const dev_caps & current_caps() {
    static dev_caps caps[64];
    static std::atomic<int> have[64];
    const int dev = current_device_slot();
    if (have[dev].load(std::memory_order_acquire) != 2) {
        int expect = 0;
        if (have[dev].compare_exchange_strong(expect, 1, std::memory_order_acq_rel)) {
            cudaDeviceGetAttribute(&caps[dev].sm_count, cudaDevAttrMultiProcessorCount, dev);
            cudaDeviceGetAttribute(&caps[dev].smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
            have[dev].store(2, std::memory_order_release);
        } else {
            while (have[dev].load(std::memory_order_acquire) != 2) { }
        }
    }
    return caps[dev];
}
// raise a kernel's dynamic shared-memory limit on the current device (monotone; safe to race: the attribute is
// only ever raised, and a thread never launches with more than it has itself asked for)
struct smem_limit {
    std::atomic<size_t> set[64];
    template <typename K> int ensure(K kern, size_t smem) {
        const int dev = current_device_slot();
        if (smem > set[dev].load(std::memory_order_acquire)) {
            const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
            if (e != cudaSuccess) return -(int) e;
            size_t cur = set[dev].load(std::memory_order_relaxed);
            while (cur < smem && !set[dev].compare_exchange_weak(cur, smem, std::memory_order_release)) { }
        }
        return B200_OK;
    }
};

// tuning knobs (ring depth, slot sizes ..): environment, read once per process at first use
inline int env_int(const char * name, int dflt) {
    const char * v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

template <int T, int NCOLS, int IPL>
static int launch_stream(stream_args a, int nw, cudaStream_t st) {
    const dev_caps & caps = current_caps();
    static const int S_env    = env_int("GGML_B200_MMVQ_STAGES", 4);
    static const int slot_env = env_int("GGML_B200_MMVQ_SLOT_BYTES", 8192);
    static const int cps_env  = env_int("GGML_B200_MMVQ_CTAS_PER_SM", 0);

    a.S = S_env < 1 ? 1 : (S_env > 8 ? 8 : S_env);
    a.R = slot_env / a.row_bytes;
    if (a.R < 1) a.R = 1;
    if (a.R > 8) a.R = 8;
    if (a.R > a.m) a.R = (int) a.m;
    a.row_pitch = (a.row_bytes + 15 + 15 + 16) & ~15;          // aligned superset + one slack word
    const int slot_bytes = a.contiguous ? (15 + a.R * a.row_bytes + 15 + 16) : a.R * a.row_pitch;
    a.slot_stride = (slot_bytes + 127) & ~127;
    const int part_bytes = 2 * nw * a.R * NCOLS * (int) sizeof(float);
    a.slots_off = (64 + part_bytes + 127) & ~127;
    const size_t smem = (size_t) a.slots_off + (size_t) a.S * a.slot_stride;
    if (smem > (size_t) caps.smem_optin) return B200_STREAM_NOT_APPLICABLE;

    auto kern = k_mmvq_stream<T, NCOLS, IPL>;
    static smem_limit lim;
    if (const int rc = lim.ensure(kern, smem)) return rc;
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, nw * 32, smem);
    if (occ < 1) return B200_STREAM_NOT_APPLICABLE;
    int cps = cps_env > 0 ? cps_env : (int) ((98304 + (size_t) a.S * a.slot_stride - 1) / ((size_t) a.S * a.slot_stride));
    if (cps < 2) cps = 2;
    if (cps > occ) cps = occ;
    const int64_t channels = a.ids ? a.dst_nb3 / a.dst_nb2 * a.n_tok_grid : 1;     // MUL_MAT_ID: one grid plane per (slot, token)
    int64_t grid = ceil_div((int64_t) caps.sm_count * cps, channels);
    const int64_t max_grid = ceil_div(a.m, a.R);
    if (grid > max_grid) grid = max_grid;
    a.rows_per_cta = (int) ceil_div(a.m, grid);
    grid = ceil_div(a.m, a.rows_per_cta);
    if (a.ids) kern<<<dim3((unsigned) grid, (unsigned) (a.dst_nb3 / a.dst_nb2), (unsigned) a.n_tok_grid), nw * 32, smem, st>>>(a);
    else       kern<<<(unsigned) grid, nw * 32, smem, st>>>(a);
    b200_count_launch(1);
    B200_RETURN_IF_LAUNCH_FAILED();
    return B200_OK;
}

template <int T, int NCOLS>
static int launch_stream_ipl(const stream_args & a, cudaStream_t st) {
    // smallest IPL whose warp count fits the register-driven thread limit
    {
        const int nw = (int) ceil_div(a.items, 32);
        if (nw * 32 <= stream_cfg<NCOLS, 1>::max_threads) return launch_stream<T, NCOLS, 1>(a, nw, st);
    }
    {
        const int nw = (int) ceil_div(a.items, 64);
        if (nw * 32 <= stream_cfg<NCOLS, 2>::max_threads) return launch_stream<T, NCOLS, 2>(a, nw, st);
    }
    return B200_STREAM_NOT_APPLICABLE;
}

template <int T>
static int launch_stream_t(stream_args a, cudaStream_t st) {
    if (!item<T>::ok(a.k, a.row_stride, reinterpret_cast<uintptr_t>(a.w))) return B200_STREAM_NOT_APPLICABLE;
    a.items = (int) (a.k / item<T>::WPI);
    if (a.n == 1) return launch_stream_ipl<T, 1>(a, st);
    if (a.n == 2) return launch_stream_ipl<T, 2>(a, st);
    if (a.n <= 4) return launch_stream_ipl<T, 4>(a, st);
    return launch_stream_ipl<T, 8>(a, st);
}

// Returns B200_OK if the streaming kernel ran, B200_STREAM_NOT_APPLICABLE if the caller should use
// the generic kernel, or an error.
// moe != NULL: MUL_MAT_ID -- one column per (slot, token) channel, the expert looked up on the device
struct stream_moe { const uint8_t * ids; int64_t ids_nb0, ids_nb1, w_nb2, n_as, ne11, n_used, n_tok; unsigned * bad; };
static int mmvq_stream_try(int type, const uint8_t * w, int64_t row_stride, int64_t m, int64_t k,
                           const int8_t * yq, const float * yd, int64_t n, float * dst, int64_t dst_col_stride,
                           cudaStream_t st, const stream_moe * moe = nullptr) {
    if (b200_option(B200_OPT_MMVQ_GENERIC)) return B200_STREAM_NOT_APPLICABLE;
    if (m < 1 || n < 1 || n > 8 || k < 256) return B200_STREAM_NOT_APPLICABLE;
    const int64_t row_bytes = k / type_qk(type) * type_bytes(type);
    if (row_bytes > 60000 || row_stride < row_bytes) return B200_STREAM_NOT_APPLICABLE;
    stream_args a;
    a.w = w; a.row_stride = row_stride; a.m = m; a.row_bytes = (int) row_bytes; a.items = 0;
    a.yq = yq; a.yd = yd; a.k = k; a.n = (int) n; a.dst = dst; a.dst_col_stride = dst_col_stride;
    a.contiguous = row_stride == row_bytes ? 1 : 0;
    a.rows_per_cta = 0; a.R = 0; a.S = 0; a.slot_stride = 0; a.slots_off = 0; a.row_pitch = 0;
    a.ids = nullptr; a.ids_nb0 = a.ids_nb1 = a.w_nb2 = a.dst_nb2 = a.dst_nb3 = 0; a.n_as = a.ne11 = 0; a.bad = nullptr; a.n_tok_grid = 0;
    if (moe) {
        if (n != 1 || (moe->w_nb2 & 15) || moe->n_used < 1 || moe->n_used > 65535 || moe->n_tok < 1 || moe->n_tok > 65535) return B200_STREAM_NOT_APPLICABLE;
        a.ids = moe->ids; a.ids_nb0 = moe->ids_nb0; a.ids_nb1 = moe->ids_nb1; a.w_nb2 = moe->w_nb2; a.dst_nb2 = m; a.dst_nb3 = m * moe->n_used;
        a.n_as = (int) moe->n_as; a.ne11 = (int) moe->ne11; a.bad = moe->bad; a.n_tok_grid = (int) moe->n_tok;
    }
    switch (type) {
        case T_Q4_0: return launch_stream_t<T_Q4_0>(a, st);
        case T_Q8_0: return launch_stream_t<T_Q8_0>(a, st);
        case T_Q4_K: return launch_stream_t<T_Q4_K>(a, st);
        case T_Q5_K: return launch_stream_t<T_Q5_K>(a, st);
        case T_Q6_K: return launch_stream_t<T_Q6_K>(a, st);
        default: return B200_ERR_TYPE;
    }
}

} // namespace b200
