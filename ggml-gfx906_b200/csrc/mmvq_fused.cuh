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

constexpr int FUSED_RB = 2;          // rows a warp carries at once (one shuffle tree per pair)

// ---------------------------------------------------------------- q8 activations of one item
// 128 consecutive activations = four q8_1 blocks: int8 values, fp16-rounded d, d * sum(q).
struct acts128 {
    int   q[4][8];
    float d[4];       // d of block j (the odd blocks pre-divided by 16 where the format asks for it)
    float sd[4];      // d * sum(q) of block j
};

// ================================================================= per-format half-block items
template <int T> struct hitem;

// ---- Q4_K: block 144 B = d,dmin | scales[12] | qs[128]; item h of block b: sub-blocks 4h..4h+3,
// quant bytes qs[64h, 64h+64): byte 32g+l holds element 64g+l (low nibble) and 64g+32+l (high nibble).
template <> struct hitem<T_Q4_K> {
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
};

// ---- Q5_K: block 176 B = d,dmin | scales[12] | qh[32] | qs[128]; bit j of qh[l] is the 5th bit of
// element l of sub-block j.
template <> struct hitem<T_Q5_K> {
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
};

} // namespace b200
