// mmvq_cols.cuh — decode at batch 2..8 (MUL_MAT with a few activation columns), Q4_K / Q5_K / Q6_K / Q4_0 / Q8_0 weights:
// ONE launch pair per list of independent nodes, int8 tensor-core MMA on the weight nibbles as they are.
//
// The reference sends n <= 8 through mul_mat_vec_q with ncols_dst template instances (src/ggml-cuda/mmvq.cu:139-226,
// one dp4a chain per column) and larger n to the MMQ tiles (mmq.cuh).  On B200 the dp4a chains make the decode
// list kernel issue-bound as soon as there is more than one column (measured: time = n x the one-column time even
// with the weights in L2), so here a warp multiplies a 16-row x 32-weight sub-block by the 8 (padded) columns
// with one mma.sync.m16n8k32 (u8 weights x s8 activations) and applies the block scales to the 16x8 int32 result:
//   * the k4 nibble layout IS the A fragment: lane (g, t) takes bytes 8t..8t+7 of a 32-byte quant chunk of rows g
//     and g + 8 (the k order inside a 32-block is free as long as A and B agree); the low nibbles (x & 0x0F) are
//     sub-block 2j, the high nibbles enter as x & 0xF0 = 16 q (u8), their int32 sums are shifted back exactly;
//   * activations are quantized once per list with the arithmetic of ggml-cpu's K-quant MUL_MAT, Q8_K
//     (quantize_row_q8_K_ref, src/ggml-quants.c:2555-2592: ONE f32 scale per 256 values and column) -- not the q8_1
//     of the n = 1 path: with one activation scale per super-block the block scales are applied in integers
//     (acc += sc * C, one IMAD per result instead of I2F + FMUL + FFMA), the -dmin * m * sum(q) term of Q4_K / Q5_K
//     and the -32 offset of Q6_K become two tiny MMAs on the per-block activation sums (bsums, split into two
//     bytes), and a super-block costs ~180 instructions per warp instead of 431 (round 1);
//   * the shipped kernel is k_mmvq_cols_ring<family> further down (one persistent CTA per SM, 32-row tiles streamed through
//     a TMA ring by three producer warps; one instantiation per format family, Q4_0 / Q8_0 and Q6_K rows at 2-byte
//     alignment included); the k_mmvq_cols<KS> of this first part (work unit = (node, 16-row tile) dealt to two CTAs per
//     SM, per-lane loads behind L2 prefetches, K-quants only) is the round-1 structure kept behind GGML_B200_COLS_RING=0;
//     in both, partial 16x8 tiles meet in shared memory and are added in fixed warp order: a row's value does not
//     depend on the grid, a column's value does not depend on the other columns.
// Values: NMSE <= 1e-9 against the oracle's Q8_K arithmetic (mode 2), which is ggml-cpu's to 1e-10.
#pragma once
#include "mmvq_batch.cuh"
#include <type_traits>

namespace b200 {

#ifndef B200_COLS_NW
#define B200_COLS_NW 8
#define B200_COLS_CTAS 2
#endif
constexpr int COLS_NW        = B200_COLS_NW;     // warps per CTA
constexpr int COLS_CTAS      = B200_COLS_CTAS;   // CTAs per SM (122 registers per thread; 3 or 4 CTAs measured 9 / 24 % slower)
constexpr int COLS_MAX_NODES = 384;    // 72 B each in the kernel parameter block (27 KB)
constexpr int COLS_N         = 8;      // columns of the MMA tile
#ifndef B200_COLS_PREFETCH_UNITS
#define B200_COLS_PREFETCH_UNITS 2
#endif
constexpr int COLS_PREFETCH_UNITS = B200_COLS_PREFETCH_UNITS;   // tiles ahead of the current one that are requested into L2

struct cols_node {
    const uint8_t * w;        // weights, rows of k4 super-blocks
    const float   * x;        // activations [n][x_stride floats]
    float         * dst;      // [n][m]
    int8_t        * yq;       // workspace: int8 [k / 32][8][32] (the 8 columns of a 32-block side by side: a B-fragment load touches 2 lines)
    uint8_t       * ym;       // workspace: per super-block f32 d [8 columns], then the activation sums (cols_meta below)
    int64_t row_stride;
    int   m, k, n, x_stride, unit0;     // unit0: index of this node's first 16-row tile in the launch
    unsigned char type, quantize, perm, unal;
    // perm (Q4_0 / Q8_0 / Q6_K nodes of the ring kernel): the int8 of a 32-block sit as 0..3 16..19 4..7 20..23 .. so that the two B
    //       words of a lane are one 8-byte load;  unal: Q6_K rows that are not 16-byte multiples (ring family 4)
};
static_assert(sizeof(cols_node) == 72, "cols_node is part of the kernel parameter block");
static_assert(COLS_NW == 8, "k_mmvq_cols<8> splits K over the 8 warps of a CTA");

struct cols_args {
    int count, units, kmax, nq;
    cols_node nodes[COLS_MAX_NODES];
    unsigned short qlist[COLS_MAX_NODES];      // the nq nodes whose activations k_quantize_cols has to produce (cols_node::quantize)
};
static_assert(sizeof(cols_args) <= 32764, "cols_args is a kernel parameter");

// The planes behind yq, for nsb = k / 256 super-blocks (block_q8_K of the reference is f32 d | i8 qs[256] | i16 bsums[16]):
//   d   f32   [nsb][8 columns]
//   s32 uint2 [nsb][8 columns][2]  word pair t: .x = low bytes (sum & 63), .y = high bytes (sum >> 6) of the sums of the
//                                  32-blocks 4t..4t+3 -- the B fragments of the m16n8k16 MMAs against the 6-bit mins
//   s16 uint2 [nsb][8 columns][4]  the same for the sums of 16 (bsums), groups 4t..4t+3 -- against Q6_K's int8 scales
//   n32 int32 [nsb][8 sub-blocks][8 columns]  -8 * (sum of the 32-block) + COLS_I2F_BIAS: the C input of Q4_0's MMAs (weights enter as
//                                  q, not q - 8); the bias makes the int32 result the bit pattern of the float 1.5 * 2^23 + result, so
//                                  the int -> float conversion (I2F: a quarter-rate instruction, 4 per block) is one FADD
constexpr int   COLS_I2F_BIAS  = 0x4B400000;       // bits of 12582912.0f = 1.5 * 2^23: exact for |result| < 2^22 (a block's dot is < 2^19)
constexpr float COLS_I2F_BIASF = 12582912.0f;
struct cols_meta {
    const float * d; const uint2 * s32; const uint2 * s16; const int * n32;
};
__device__ __host__ __forceinline__ cols_meta cols_meta_of(const uint8_t * ym, int k) {
    const size_t nsb = (size_t) (k >> 8);
    return { reinterpret_cast<const float *>(ym), reinterpret_cast<const uint2 *>(ym + nsb * 32), reinterpret_cast<const uint2 *>(ym + nsb * 160),
             reinterpret_cast<const int *>(ym + nsb * 416) };
}

// ---------------------------------------------------------------- activations of the list: f32 -> Q8_K columns
// grid (ceil(kmax / 256 / 8), nq, 8), block 256: one WARP per super-block of one column (8 consecutive values per lane), only the
// nodes of qlist (those that do not share an earlier node's activations).  One CTA per super-block measured ~45 us per Llama-2-7B
// pass at 8 columns -- 77 000 CTAs of 1 KB, most of them empty -- against 670 us for the weights.
// Bit-exact quantize_row_q8_K_ref: the FIRST value of largest magnitude maps to -127, round to nearest even, min(127, .).
__global__ void __launch_bounds__(256) k_quantize_cols(const __grid_constant__ cols_args a) {
    const cols_node & nd = a.nodes[a.qlist[blockIdx.y]];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sb = blockIdx.x * 8 + warp, c = blockIdx.z;
    if (sb * 256 >= nd.k) return;
    const cols_meta M = cols_meta_of(nd.ym, nd.k);
    float  * dp  = const_cast<float *>(M.d) + sb * COLS_N + c;
    uint8_t * p32 = reinterpret_cast<uint8_t *>(const_cast<uint2 *>(M.s32) + (sb * COLS_N + c) * 2);
    uint8_t * p16 = reinterpret_cast<uint8_t *>(const_cast<uint2 *>(M.s16) + (sb * COLS_N + c) * 4);
    int * pn = const_cast<int *>(M.n32) + (size_t) sb * 8 * COLS_N + c;         // + sub-block * COLS_N
    if (c >= nd.n) {                                      // padding columns: zero scale and sums, q never read as non-zero
        if (lane == 0) *dp = 0.0f;
        if (lane < 4)  reinterpret_cast<uint32_t *>(p32)[lane] = 0u;
        if (lane < 8)  { reinterpret_cast<uint32_t *>(p16)[lane] = 0u; pn[lane * COLS_N] = COLS_I2F_BIAS; }
        return;
    }
    const float * xp = nd.x + (size_t) c * nd.x_stride + sb * 256 + lane * 8;
    float x[8];
    if ((reinterpret_cast<uintptr_t>(xp) & 15) == 0) {    // the same answer for every lane
        const float4 v0 = __ldg(reinterpret_cast<const float4 *>(xp)), v1 = __ldg(reinterpret_cast<const float4 *>(xp) + 1);
        x[0] = v0.x; x[1] = v0.y; x[2] = v0.z; x[3] = v0.w; x[4] = v1.x; x[5] = v1.y; x[6] = v1.z; x[7] = v1.w;
    } else {
#pragma unroll
        for (int e = 0; e < 8; e++) x[e] = __ldg(xp + e);
    }
    // argmax |x| with the lowest index winning ties: key = |x| bits : (255 - j), j = 8 lane + e
    unsigned long long key = 0ull;
#pragma unroll
    for (int e = 0; e < 8; e++) {
        const unsigned long long ke = ((unsigned long long) (__float_as_uint(x[e]) & 0x7FFFFFFFu) << 8) | (unsigned) (255 - (lane * 8 + e));
        key = ke > key ? ke : key;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) { const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, key, off); key = o > key ? o : key; }
    const int jmax = 255 - (int) (key & 0xFFu);
    float xs = x[0];
#pragma unroll
    for (int e = 1; e < 8; e++) xs = (jmax & 7) == e ? x[e] : xs;
    const float mx = __shfl_sync(0xFFFFFFFFu, xs, jmax >> 3);
    int q[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    float d = 0.0f;
    if ((key >> 8) != 0ull) {                             // amax != 0
        const float iscale = __fdiv_rn(-127.0f, mx);
#pragma unroll
        for (int e = 0; e < 8; e++) q[e] = min(127, __float2int_rn(__fmul_rn(iscale, x[e])));
        d = __fdiv_rn(1.0f, iscale);
    }
    uint2 pk;
    pk.x = (uint32_t) (q[0] & 0xFF) | ((uint32_t) (q[1] & 0xFF) << 8) | ((uint32_t) (q[2] & 0xFF) << 16) | ((uint32_t) (q[3] & 0xFF) << 24);
    pk.y = (uint32_t) (q[4] & 0xFF) | ((uint32_t) (q[5] & 0xFF) << 8) | ((uint32_t) (q[6] & 0xFF) << 16) | ((uint32_t) (q[7] & 0xFF) << 24);
    // 32-block lane / 4, bytes 8 (lane & 3)..
    int8_t * yb = nd.yq + ((size_t) (sb * 8 + (lane >> 2)) * COLS_N + c) * 32;
    if (!nd.perm) *reinterpret_cast<uint2 *>(yb + (lane & 3) * 8) = pk;
    else {
        // elements j0 = 8 (lane & 3) .. j0 + 3 and j0 + 4 .. j0 + 7; four elements from j go to byte 2 j (j < 16) or 2 (j - 16) + 4
        const int j0 = (lane & 3) * 8, p0 = j0 < 16 ? 2 * j0 : 2 * (j0 - 16) + 4;
        *reinterpret_cast<uint32_t *>(yb + p0) = pk.x; *reinterpret_cast<uint32_t *>(yb + p0 + 8) = pk.y;
    }
    int sum = q[0] + q[1] + q[2] + q[3] + q[4] + q[5] + q[6] + q[7];
    sum += __shfl_xor_sync(0xFFFFFFFFu, sum, 1);
    if ((lane & 1) == 0) {                                // sums of 16: group lane / 2
        const int gi = lane >> 1;
        p16[(gi >> 2) * 8 + (gi & 3)]     = (uint8_t) (sum & 63);
        p16[(gi >> 2) * 8 + 4 + (gi & 3)] = (uint8_t) (sum >> 6);
    }
    sum += __shfl_xor_sync(0xFFFFFFFFu, sum, 2);
    if ((lane & 3) == 0) {                                // sum of the 32-block lane / 4
        const int bi = lane >> 2;
        p32[(bi >> 2) * 8 + (bi & 3)]     = (uint8_t) (sum & 63);
        p32[(bi >> 2) * 8 + 4 + (bi & 3)] = (uint8_t) (sum >> 6);
        pn[bi * COLS_N] = COLS_I2F_BIAS - 8 * sum;
    }
    if (lane == 0) *dp = d;
}

// ---------------------------------------------------------------- the MMAs (C = 0: the scales differ per 32-block)
// m16n8k32, A = u8 weights (rows g / g + 8, k slots 4t.. and 16 + 4t..), B = s8 activations (column g)
__device__ __forceinline__ void mma_u8s8_16x8x32(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%10, %10, %10, %10};"
        : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
}
// the same with a C input (Q4_0: -8 * sum of the activations) and with signed weights (Q8_0)
__device__ __forceinline__ void mma_u8s8_16x8x32_c(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1, int c01, int c11) {
    asm("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%10, %11, %10, %11};"
        : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(c01), "r"(c11));
}
__device__ __forceinline__ void mma_s8s8_16x8x32(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1, int c) {
    asm("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%10, %10, %10, %10};"
        : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(c));
}
// m16n8k16, A = u8 (mins, 6-bit weights) or s8 (Q6_K scales), B = s8
template <bool SIGNED_A>
__device__ __forceinline__ void mma_16x8x16(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    if (SIGNED_A)
        asm("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%7, %7, %7, %7};"
            : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]) : "r"(a0), "r"(a1), "r"(b0), "r"(0));
    else
        asm("mma.sync.aligned.m16n8k16.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%7, %7, %7, %7};"
            : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]) : "r"(a0), "r"(a1), "r"(b0), "r"(0));
}
// byte i of w, zero- / sign-extended (one PRMT each; __byte_perm drops prmt's sign-replicate selector bit)
template <int I> __device__ __forceinline__ int byte_u(uint32_t w) { return (int) __byte_perm(w, 0u, 0x4440 + I); }
template <int I> __device__ __forceinline__ int byte_s(uint32_t w) {
    int r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(0), "n"(I | ((I | 8) << 4) | ((I | 8) << 8) | ((I | 8) << 12)));
    return r;
}

// streaming loads of the weights: read-only path, not kept in L1 (L1 is for the activation columns)
__device__ __forceinline__ uint2 ldg_stream64(const void * p) {
    uint2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ uint4 ldg_stream128(const void * p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// bytes of 256 weights of a row (one super-block of the K-quants, eight blocks of Q4_0 / Q8_0)
__device__ __host__ __forceinline__ int cols_block_bytes(int type) { return type == T_Q6_K ? 210 : (type == T_Q5_K ? 176 : (type == T_Q8_0 ? 272 : 144)); }

// HBM -> L2 prefetch of a whole weight row (16-byte aligned, a multiple of 16 bytes): no registers, no completion
__device__ __forceinline__ void l2_prefetch_bulk(const void * p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p), "r"(bytes) : "memory");
}

// The words of one 256-weight super-block of one row that lane (g, t) needs.  The MMA's k order inside a 32-block is
// free as long as A and B agree: lane t takes bytes 8t..8t+7 of the 32-byte quant chunk (one 8-byte load) as its k
// slots 4t..4t+3 and 16+4t..16+4t+3, and the same bytes of the activation block.
struct cols_item {
    uint4 hdr;                // d, dmin (f16 x 2) + 12 scale bytes
    uint2 q[4];               // [32-byte quant chunk] bytes 8t..8t+7
    uint2 h;                  // Q5_K: high-bit bytes 8t..8t+7
};

__device__ __forceinline__ void cols_load_any(cols_item & it, const uint8_t * blk, int t, bool q5) {
    it.hdr = ldg_stream128(blk);
    const uint8_t * qs = blk + (q5 ? 48 : 16) + 8 * t;
#pragma unroll
    for (int p = 0; p < 4; p++) it.q[p] = ldg_stream64(qs + 32 * p);
    if (q5) it.h = ldg_stream64(blk + 16 + 8 * t);
}

// bit S of every byte of h, moved to bit 4
template <int S> __device__ __forceinline__ uint32_t cols_hbit(uint32_t h) {
    return (S <= 4 ? (h << (4 - S)) : (h >> (S - 4))) & 0x10101010u;
}

// One super-block of rows g and g + 8 against the 8 columns: 8 MMAs (one per 32-weight sub-block), the 6-bit block
// scales applied to the int32 results, the mins through two small MMAs on the activations' block sums.
//   sum_k w y = d dy sum_b sc_b (q_b . y_b) - dmin dy sum_b m_b sum(y_b)
// While it runs it rolls the NEXT super-block of the warp (blocks at na / nb, of a Q4_K or Q5_K row: nq5) into ia / ib,
// each word as soon as its last use is behind: the loads stay in flight for one whole super-block of work (an L2 round
// trip) without a second set of registers.  na == nullptr: nothing follows.
template <bool Q5>
__device__ __forceinline__ void cols_sb_mma(cols_item & ia, cols_item & ib, const uint2 (&b)[8], const float2 dy,
                                            const uint2 s32, const int t, float (&acc)[4],
                                            const uint8_t * na, const uint8_t * nb, const bool nq5) {
    // get_scale_min_k4, four sub-blocks per word
    const uint32_t scA[2] = { ia.hdr.y & 0x3F3F3F3Fu, (ia.hdr.w & 0x0F0F0F0Fu) | ((ia.hdr.y >> 2) & 0x30303030u) };
    const uint32_t mnA[2] = { ia.hdr.z & 0x3F3F3F3Fu, ((ia.hdr.w >> 4) & 0x0F0F0F0Fu) | ((ia.hdr.z >> 2) & 0x30303030u) };
    const uint32_t scB[2] = { ib.hdr.y & 0x3F3F3F3Fu, (ib.hdr.w & 0x0F0F0F0Fu) | ((ib.hdr.y >> 2) & 0x30303030u) };
    const uint32_t mnB[2] = { ib.hdr.z & 0x3F3F3F3Fu, ((ib.hdr.w >> 4) & 0x0F0F0F0Fu) | ((ib.hdr.z >> 2) & 0x30303030u) };
    const float2 dmA = __half22float2(*reinterpret_cast<const __half2 *>(&ia.hdr.x));
    const float2 dmB = __half22float2(*reinterpret_cast<const __half2 *>(&ib.hdr.x));
    const bool more = na != nullptr;
    const uint8_t * nqa = na + (nq5 ? 48 : 16) + 8 * t, * nqb = nb + (nq5 ? 48 : 16) + 8 * t;
    if (more) { ia.hdr = ldg_stream128(na); ib.hdr = ldg_stream128(nb); }
    int lo[4] = { 0, 0, 0, 0 }, hi[4] = { 0, 0, 0, 0 };      // hi: the high-nibble sub-blocks of Q4_K, 16 x too large
    auto one = [&](auto Q) {
        constexpr int q = decltype(Q)::value, p = q >> 1;
        uint32_t a0, a1, a2, a3;
        if (!Q5) {
            constexpr uint32_t msk = (q & 1) ? 0xF0F0F0F0u : 0x0F0F0F0Fu;
            a0 = ia.q[p].x & msk; a1 = ib.q[p].x & msk; a2 = ia.q[p].y & msk; a3 = ib.q[p].y & msk;
        } else {
            constexpr int sh = (q & 1) * 4;                    // bit q of qh belongs to sub-block q
            a0 = ((ia.q[p].x >> sh) & 0x0F0F0F0Fu) | cols_hbit<q>(ia.h.x); a1 = ((ib.q[p].x >> sh) & 0x0F0F0F0Fu) | cols_hbit<q>(ib.h.x);
            a2 = ((ia.q[p].y >> sh) & 0x0F0F0F0Fu) | cols_hbit<q>(ia.h.y); a3 = ((ib.q[p].y >> sh) & 0x0F0F0F0Fu) | cols_hbit<q>(ib.h.y);
        }
        int c[4];
        mma_u8s8_16x8x32(c, a0, a1, a2, a3, b[q].x, b[q].y);
        const int sA = byte_u<q & 3>(scA[q >> 2]), sB = byte_u<q & 3>(scB[q >> 2]);
        if (!Q5 && (q & 1)) { hi[0] += sA * c[0]; hi[1] += sA * c[1]; hi[2] += sB * c[2]; hi[3] += sB * c[3]; }
        else                { lo[0] += sA * c[0]; lo[1] += sA * c[1]; lo[2] += sB * c[2]; lo[3] += sB * c[3]; }
        if ((q & 1) && more) { ia.q[p] = ldg_stream64(nqa + 32 * p); ib.q[p] = ldg_stream64(nqb + 32 * p); }     // chunk p is done
    };
    one(std::integral_constant<int, 0>()); one(std::integral_constant<int, 1>()); one(std::integral_constant<int, 2>());
    one(std::integral_constant<int, 3>()); one(std::integral_constant<int, 4>()); one(std::integral_constant<int, 5>());
    one(std::integral_constant<int, 6>()); one(std::integral_constant<int, 7>());
    // sum_b m_b * sum32_b: A = the mins (k = sub-block; lanes t >= 2 hold zero B words), sums = lo + 64 hi
    int cl[4], ch[4];
    const uint32_t am0 = t == 0 ? mnA[0] : mnA[1], am1 = t == 0 ? mnB[0] : mnB[1];
    mma_16x8x16<false>(cl, am0, am1, s32.x);
    mma_16x8x16<false>(ch, am0, am1, s32.y);
    if (more && nq5) { ia.h = ldg_stream64(na + 16 + 8 * t); ib.h = ldg_stream64(nb + 16 + 8 * t); }
    acc[0] += dy.x * (dmA.x * (float) (lo[0] + (hi[0] >> 4)) - dmA.y * (float) (cl[0] + 64 * ch[0]));
    acc[1] += dy.y * (dmA.x * (float) (lo[1] + (hi[1] >> 4)) - dmA.y * (float) (cl[1] + 64 * ch[1]));
    acc[2] += dy.x * (dmB.x * (float) (lo[2] + (hi[2] >> 4)) - dmB.y * (float) (cl[2] + 64 * ch[2]));
    acc[3] += dy.y * (dmB.x * (float) (lo[3] + (hi[3] >> 4)) - dmB.y * (float) (cl[3] + 64 * ch[3]));
}

// ---------------------------------------------------------------- Q6_K: one m16n8k16 MMA per 16-weight scale group
// block = ql[128] qh[64] sc[16] (int8) d (f16) = 210 bytes: blocks sit at 2-byte alignment.  With w and the row
// stride multiples of 4 the blocks of even index are 4-byte aligned and those of odd index are 2 bytes off, the
// same for every lane: odd blocks read the two aligned words around each word and funnel-shift.
// Lane (g, t) takes bytes 4t..4t+3 of each 16-byte piece: for half h, piece s, word j of ql serves the groups
// r = j (low nibbles) and r = j + 2 (high nibbles); the qh word serves r = 0..3 (2 bits each).
//   sum_k w y = d dy (sum_g sc_g (q_g . y_g) - 32 sum_g sc_g sum16_g),  q in 0..63 (u8), sc int8
__device__ __forceinline__ uint32_t ldg_stream32(const void * p) {
    uint32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
struct cols_item6 {
    uint32_t ql[2][2][2];     // [half h][word j: l / l + 32][piece s]
    uint32_t qh[2][2];        // [half h][piece s]
    uint32_t sc[4];           // 16 int8 scales
    uint32_t d;               // f16 bits
};
template <bool SMEM> __device__ __forceinline__ uint32_t cols_ld32(const uint8_t * p) {
    if (SMEM) return *reinterpret_cast<const uint32_t *>(p);
    return ldg_stream32(p);
}
template <bool ODD, bool SMEM>
__device__ __forceinline__ uint32_t cols_ld6(const uint8_t * blk, int off) {
    if (!ODD) return cols_ld32<SMEM>(blk + off);
    return __funnelshift_r(cols_ld32<SMEM>(blk + off - 2), cols_ld32<SMEM>(blk + off + 2), 16);
}
template <bool ODD, bool SMEM = false>
__device__ __forceinline__ void cols_load6(cols_item6 & it, const uint8_t * blk, int t) {
#pragma unroll
    for (int h = 0; h < 2; h++)
#pragma unroll
        for (int s = 0; s < 2; s++) {
            it.ql[h][0][s] = cols_ld6<ODD, SMEM>(blk, 64 * h + 16 * s + 4 * t);
            it.ql[h][1][s] = cols_ld6<ODD, SMEM>(blk, 64 * h + 32 + 16 * s + 4 * t);
            it.qh[h][s]    = cols_ld6<ODD, SMEM>(blk, 128 + 32 * h + 16 * s + 4 * t);
        }
    if (!ODD) {
#pragma unroll
        for (int i = 0; i < 4; i++) it.sc[i] = cols_ld32<SMEM>(blk + 192 + 4 * i);
    } else {
        uint32_t w[5];
#pragma unroll
        for (int i = 0; i < 5; i++) w[i] = cols_ld32<SMEM>(blk + 190 + 4 * i);
#pragma unroll
        for (int i = 0; i < 4; i++) it.sc[i] = __funnelshift_r(w[i], w[i + 1], 16);
    }
    it.d = *reinterpret_cast<const unsigned short *>(blk + 208);
}
// shared memory, block at any 2-byte alignment (known at run time only): every word is two aligned loads and a funnel shift
__device__ __forceinline__ void cols_load6_var(cols_item6 & it, const uint8_t * blk, int t) {
    const uint8_t * b4 = reinterpret_cast<const uint8_t *>(reinterpret_cast<uintptr_t>(blk) & ~(uintptr_t) 3);
    const uint32_t sh = (uint32_t) (reinterpret_cast<uintptr_t>(blk) & 2) * 8;
    auto ld = [&](int off) { return __funnelshift_r(*reinterpret_cast<const uint32_t *>(b4 + off), *reinterpret_cast<const uint32_t *>(b4 + off + 4), sh); };
#pragma unroll
    for (int h = 0; h < 2; h++)
#pragma unroll
        for (int s = 0; s < 2; s++) {
            it.ql[h][0][s] = ld(64 * h + 16 * s + 4 * t);
            it.ql[h][1][s] = ld(64 * h + 32 + 16 * s + 4 * t);
            it.qh[h][s]    = ld(128 + 32 * h + 16 * s + 4 * t);
        }
    uint32_t w[5];
#pragma unroll
    for (int i = 0; i < 5; i++) w[i] = *reinterpret_cast<const uint32_t *>(b4 + 192 + 4 * i);
#pragma unroll
    for (int i = 0; i < 4; i++) it.sc[i] = __funnelshift_r(w[i], w[i + 1], sh);
    it.d = *reinterpret_cast<const unsigned short *>(blk + 208);
}
// yq_col: the column's int8 plane, [32-block][8 columns][32] from block 8 * item on (global memory, or the stage of the ring)
// PERM (ring kernel): the int8 of a 32-block were stored as 0..3 16..19 4..7 20..23 .. (cols_node::perm), so the B words of the
// two 16-weight groups of a block are one 8-byte load at 8 t (the 4-byte loads 16 bytes apart collided between row groups g, g + 4)
template <bool SMEM = false, bool PERM = false>
__device__ __forceinline__ void cols_item6_mma(const cols_item6 & ia, const cols_item6 & ib, int item, int t,
                                               const int8_t * yq_col, const float2 dy, const uint2 s16, bool col_ok, float (&acc)[4]) {
    int ai[4] = { 0, 0, 0, 0 };
    uint2 bb = make_uint2(0u, 0u);
    auto group = [&](auto H, auto R, auto S) {
        constexpr int h = decltype(H)::value, r = decltype(R)::value, s = decltype(S)::value;
        const int B = item * 8 + 4 * h + r;                         // the 32-block of the activations
        if (PERM && s == 0) bb = *reinterpret_cast<const uint2 *>(yq_col + B * (32 * COLS_N) + 8 * t);
        const uint32_t a0 = ((ia.ql[h][r & 1][s] >> (4 * (r >> 1))) & 0x0F0F0F0Fu) | (((ia.qh[h][s] >> (2 * r)) & 0x03030303u) << 4);
        const uint32_t a1 = ((ib.ql[h][r & 1][s] >> (4 * (r >> 1))) & 0x0F0F0F0Fu) | (((ib.qh[h][s] >> (2 * r)) & 0x03030303u) << 4);
        uint32_t b0 = 0u;
        if (PERM) b0 = s ? bb.y : bb.x;
        else if (SMEM) b0 = *reinterpret_cast<const uint32_t *>(yq_col + B * (32 * COLS_N) + 16 * s + 4 * t);
        else if (col_ok) b0 = __ldg(reinterpret_cast<const uint32_t *>(yq_col + (size_t) B * (32 * COLS_N) + 16 * s + 4 * t));
        int c[4];
        mma_16x8x16<false>(c, a0, a1, b0);
        constexpr int si = 8 * h + 2 * r + s;                       // scale of this 16-weight group
        const int fA = byte_s<si & 3>(ia.sc[si >> 2]), fB = byte_s<si & 3>(ib.sc[si >> 2]);
        ai[0] += fA * c[0]; ai[1] += fA * c[1]; ai[2] += fB * c[2]; ai[3] += fB * c[3];
    };
    using I0 = std::integral_constant<int, 0>; using I1 = std::integral_constant<int, 1>;
    using I2 = std::integral_constant<int, 2>; using I3 = std::integral_constant<int, 3>;
    group(I0(), I0(), I0()); group(I0(), I0(), I1()); group(I0(), I1(), I0()); group(I0(), I1(), I1());
    group(I0(), I2(), I0()); group(I0(), I2(), I1()); group(I0(), I3(), I0()); group(I0(), I3(), I1());
    group(I1(), I0(), I0()); group(I1(), I0(), I1()); group(I1(), I1(), I0()); group(I1(), I1(), I1());
    group(I1(), I2(), I0()); group(I1(), I2(), I1()); group(I1(), I3(), I0()); group(I1(), I3(), I1());
    // sum_g sc_g * sum16_g: A = the int8 scales (k = group: bytes 4t..4t+3), sums = lo + 64 hi
    const uint32_t sa = t == 0 ? ia.sc[0] : (t == 1 ? ia.sc[1] : (t == 2 ? ia.sc[2] : ia.sc[3]));
    const uint32_t sb = t == 0 ? ib.sc[0] : (t == 1 ? ib.sc[1] : (t == 2 ? ib.sc[2] : ib.sc[3]));
    int cl[4], ch[4];
    mma_16x8x16<true>(cl, sa, sb, s16.x);
    mma_16x8x16<true>(ch, sa, sb, s16.y);
    const float dA = f16_bits_to_f32(ia.d), dB = f16_bits_to_f32(ib.d);
    acc[0] += (dA * dy.x) * (float) (ai[0] - 32 * (cl[0] + 64 * ch[0]));
    acc[1] += (dA * dy.y) * (float) (ai[1] - 32 * (cl[1] + 64 * ch[1]));
    acc[2] += (dB * dy.x) * (float) (ai[2] - 32 * (cl[2] + 64 * ch[2]));
    acc[3] += (dB * dy.y) * (float) (ai[3] - 32 * (cl[3] + 64 * ch[3]));
}

// KS = 8 (default): the 8 warps of a CTA split K of one tile, partial tiles meet in shared memory;
// KS = 1 (tuning variant): every warp owns whole tiles -- no barrier, no shared-memory reduction.
// A row's value does not depend on the grid in either; the two differ in the f32 summation order over K.
template <int KS>
__global__ void __launch_bounds__(COLS_NW * 32, COLS_CTAS) k_mmvq_cols(const __grid_constant__ cols_args a) {
    __shared__ float red[KS > 1 ? 2 : 1][KS > 1 ? COLS_NW : 1][KS > 1 ? 128 : 1];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    int ni = 0, par = 0;
    const int ustride = KS > 1 ? (int) gridDim.x : (int) gridDim.x * COLS_NW;
    const int kw = KS > 1 ? warp : 0;                       // this warp's first super-block of a tile, and its stride
    const int ptid = KS > 1 ? (int) threadIdx.x : lane;     // the thread's index among those that prefetch a tile
    int u = KS > 1 ? (int) blockIdx.x : (int) blockIdx.x * COLS_NW + warp;
    // Weights reach the SM in two steps: whole tiles are requested into L2 a few units ahead by bulk prefetches
    // (one weight row per thread 0..15; no registers, ~150 KB in flight per SM), and a warp's loads of its
    // super-block then hit L2.  The registers go to the activation fragments instead, which are requested half a
    // super-block ahead of their MMAs (ncu: the long-scoreboard stalls sat on the MMAs' B operands).
    auto prefetch_unit = [&](int up, int & nip) {
        if (up >= a.units) return;
        while (nip + 1 < a.count && up >= a.nodes[nip + 1].unit0) nip++;
        const cols_node & np = a.nodes[nip];
        const int row = (up - np.unit0) * 16 + ptid;
        if (ptid < 16 && row < np.m)
            l2_prefetch_bulk(np.w + (size_t) row * np.row_stride, (uint32_t) (((np.k >> 8) * cols_block_bytes(np.type)) & ~15));
    };
    // KS = 1: 16 warps x whole tiles ahead would not fit L2; a warp requests 8 super-blocks (1152 B of each of its 16
    // rows) at a time, 8 ahead of the one it works on, across tile boundaries
    auto prefetch_chunk = [&](const cols_node & np, int row0p, int b0) {
        const int nb = min(8, (np.k >> 8) - b0), bs = cols_block_bytes(np.type);
        if (np.type != T_Q6_K && lane < 16 && row0p + lane < np.m && nb > 0)
            l2_prefetch_bulk(np.w + (size_t) (row0p + lane) * np.row_stride + (size_t) b0 * bs, (uint32_t) (nb * bs));
    };
    const int PD = COLS_PREFETCH_UNITS;
    int nip = 0;
    if (KS > 1) { for (int d = 0; d < PD; d++) prefetch_unit(u + d * ustride, nip); }
    else if (u < a.units) {
        while (nip + 1 < a.count && u >= a.nodes[nip + 1].unit0) nip++;
        prefetch_chunk(a.nodes[nip], (u - a.nodes[nip].unit0) * 16, 0);
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");        // the quantizer of this list has finished
    cols_item ia, ib;          // the warp's current Q4_K / Q5_K super-block of rows g and g + 8 ...
    bool have = false;         // ... already requested by the step before (also across a tile boundary)
    for (; u < a.units; u += ustride, par ^= 1) {
        while (ni + 1 < a.count && u >= a.nodes[ni + 1].unit0) ni++;
        const cols_node & nd = a.nodes[ni];
        const int row0 = (u - nd.unit0) * 16, items = nd.k >> 8;
        const int ra = min(row0 + g, nd.m - 1), rb = min(row0 + g + 8, nd.m - 1);      // tail rows: recomputed, never stored
        const uint8_t * pa = nd.w + (size_t) ra * nd.row_stride;
        const uint8_t * pb = nd.w + (size_t) rb * nd.row_stride;
        if (KS > 1) prefetch_unit(u + PD * ustride, nip);
        else if (u + ustride < a.units) { while (nip + 1 < a.count && u + ustride >= a.nodes[nip + 1].unit0) nip++; }   // nip: the next tile's node
        // the warp's first super-block of its NEXT tile, if that is a Q4_K / Q5_K one
        const uint8_t * pan = nullptr, * pbn = nullptr;
        bool q5n = false;
        if (u + ustride < a.units) {
            int nin = ni;
            while (nin + 1 < a.count && u + ustride >= a.nodes[nin + 1].unit0) nin++;
            const cols_node & nn = a.nodes[nin];
            if (nn.type != T_Q6_K && (nn.k >> 8) > kw) {
                const int r0n = (u + ustride - nn.unit0) * 16;
                q5n = nn.type == T_Q5_K;
                pan = nn.w + (size_t) min(r0n + g, nn.m - 1) * nn.row_stride + (size_t) kw * (q5n ? 176 : 144);
                pbn = nn.w + (size_t) min(r0n + g + 8, nn.m - 1) * nn.row_stride + (size_t) kw * (q5n ? 176 : 144);
            }
        }
        const int8_t * yq_col = nd.yq + g * 32;
        const cols_meta M = cols_meta_of(nd.ym, nd.k);
        const bool col_ok = g < nd.n;
        float acc[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
        if (nd.type == T_Q6_K) {
            for (int it = kw; it < items; it += KS) {
                cols_item6 ia6, ib6;
                if (it & 1) { cols_load6<true>(ia6, pa + (size_t) it * 210, t);  cols_load6<true>(ib6, pb + (size_t) it * 210, t); }
                else        { cols_load6<false>(ia6, pa + (size_t) it * 210, t); cols_load6<false>(ib6, pb + (size_t) it * 210, t); }
                const float2 dy = __ldg(reinterpret_cast<const float2 *>(M.d + it * COLS_N + 2 * t));       // columns 2t, 2t + 1
                const uint2 s16 = __ldg(M.s16 + (it * COLS_N + g) * 4 + t);                                 // column g, groups 4t..4t+3
                cols_item6_mma(ia6, ib6, it, t, yq_col, dy, s16, col_ok, acc);
            }
            if (pan) { cols_load_any(ia, pan, t, q5n); cols_load_any(ib, pbn, t, q5n); }
            have = pan != nullptr;
        } else if (kw < items) {
            auto tile = [&](auto Q5) {
                constexpr bool q5 = decltype(Q5)::value;
                constexpr int bs = q5 ? 176 : 144;
                if (!have) { cols_load_any(ia, pa + (size_t) kw * bs, t, q5); cols_load_any(ib, pb + (size_t) kw * bs, t, q5); }
                for (int it = kw; it < items; it += KS) {
                    if (KS == 1 && (it & 7) == 0) {
                        if (it + 8 < items) prefetch_chunk(nd, row0, it + 8);
                        else if (u + ustride < a.units) prefetch_chunk(a.nodes[nip], (u + ustride - a.nodes[nip].unit0) * 16, 0);
                    }
                    const bool last = it + KS >= items;
                    const uint8_t * na = last ? pan : pa + (size_t) (it + KS) * bs;
                    const uint8_t * nb = last ? pbn : pb + (size_t) (it + KS) * bs;
                    uint2 b[8];
                    const int8_t * yq = yq_col + (size_t) it * (8 * 32 * COLS_N) + 8 * t;
#pragma unroll
                    for (int q = 0; q < 8; q++) {
                        b[q] = make_uint2(0u, 0u);
                        if (col_ok) b[q] = __ldg(reinterpret_cast<const uint2 *>(yq + q * (32 * COLS_N)));
                    }
                    const float2 dy = __ldg(reinterpret_cast<const float2 *>(M.d + it * COLS_N + 2 * t));   // columns 2t, 2t + 1
                    uint2 s32 = make_uint2(0u, 0u);
                    if (t < 2) s32 = __ldg(M.s32 + (it * COLS_N + g) * 2 + t);                              // column g, sub-blocks 4t..4t+3
                    cols_sb_mma<q5>(ia, ib, b, dy, s32, t, acc, na, nb, last ? q5n : q5);
                }
            };
            if (nd.type == T_Q5_K) tile(std::true_type()); else tile(std::false_type());
            have = pan != nullptr;
        } else {
            have = false;
        }
        if (KS == 1) {
            // c fragment layout: acc[0..1] = row g, columns 2t, 2t + 1; acc[2..3] = row g + 8
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int row = row0 + g + 8 * (i >> 1), col = 2 * t + (i & 1);
                if (row < nd.m && col < nd.n) nd.dst[(size_t) col * nd.m + row] = acc[i];
            }
            continue;
        }
        *reinterpret_cast<float4 *>(&red[par][warp][lane * 4]) = make_float4(acc[0], acc[1], acc[2], acc[3]);
        __syncthreads();
        // red is double-buffered by unit parity: the buffer written two units later is only touched by threads
        // that have passed the next unit's barrier, which these readers reach after their reads
        if (threadIdx.x < 128) {
            const int e = threadIdx.x, l = e >> 2, i = e & 3;
            const int row = row0 + (l >> 2) + 8 * (i >> 1), col = 2 * (l & 3) + (i & 1);
            float v = 0.0f;
#pragma unroll
            for (int w = 0; w < COLS_NW; w++) v += red[par][w][e];
            if (row < nd.m && col < nd.n) nd.dst[(size_t) col * nd.m + row] = v;
        }
    }
}

// ---------------------------------------------------------------- Q4_K / Q5_K: tiles staged through a TMA ring
// k_mmvq_cols above reads its weights with per-lane loads out of L2 (behind bulk L2 prefetches): at most ~40 KB in flight
// per SM, every 8-byte load touching eight 128-byte lines, and the kernel stood at 50 % of the HBM peak with every
// pipe below 60 % (profiles/r03c_ncu_k_mmvq_cols_q8k_summary.txt; rolling the next super-block's words into the
// registers or dropping the per-tile barrier changed nothing).  Here, as in the one-column list kernel (mmvq_batch.cuh):
// one persistent CTA per SM; a producer warp streams (32-row tile) x (8 super-blocks) chunks -- the weights and the matching
// slice of the quantized activations -- into a ring of three 62 KB stages with 1-D bulk copies (one per row + three;
// full / empty mbarriers), so ~190 KB are in flight per SM and no consumer load ever leaves the SM; sixteen consumer
// warps take one super-block of 16 rows each out of shared memory (row pitch = 32 mod 128 bytes: the 8-byte fragment
// loads of a warp cover all banks twice, the minimum); their partial 16x8 tiles go to a double-buffered shared array
// and an epilogue warp adds them in fixed order and stores -- no CTA-wide barrier anywhere.
constexpr int RING_NWC      = 16;                  // consumer warps: warp w -> rows 16 (w >> 3).., super-block (w & 7) of a chunk
constexpr int RING_NPROD    = 3;                   // producer warps: a bulk copy is a per-lane serial issue (~100 cycles each); one
                                                   // producer warp could not feed the ring (consumers waited 43 % of the time)
constexpr int RING_THREADS  = (RING_NWC + RING_NPROD + 1) * 32;   // consumers + producers + epilogue warp (20 warps: 96 registers)
constexpr int RING_CI       = 8;                   // super-blocks per chunk
constexpr int RING_ROWS     = 32;
constexpr int RING_STAGES   = 3;
// shared-memory geometry of a family: 0 = Q4_K / Q5_K (rows of 8 x 144 / 176 bytes + 32: pitch = 32 mod 128, the 8-byte
// fragment loads of the 8 row groups cover every bank twice), 1 = Q6_K (8 x 210 = 1680 = 16 mod 128: its 4-byte loads, 16
// bytes per row group, cover every bank once).  A stage = the chunk's weights, then the matching slice of the activations:
//   int8 [8 super-blocks][8 sub-blocks][8 columns][32] | f32 d [8][8 columns] | sums of 32 [8][8][2] uint2 or sums of 16 [8][8][4] uint2
// Family 2 = Q8_0 (eight 34-byte blocks = 272 bytes per 256 weights, rows of 8 x 272 + 16 = 2192 = 16 mod 128; two stages);
// family 3 = Q4_0 (eight 18-byte blocks = 144 bytes, rows of 8 x 144 + 16 = 1168 = 16 mod 128, the n32 plane in place of s32).  Its own
// instantiation: folded into family 0 behind a run-time type test it cost Q4_K 7 % (0.720 -> 0.774 ms per Llama-2-7B pass at n = 8).
template <int FAM> struct ring_geo {
    static constexpr int stages    = FAM == 2 ? 2 : RING_STAGES;
    static constexpr int pitch_max = FAM == 1 ? RING_CI * 210 : (FAM == 4 ? RING_CI * 210 + 32 : (FAM == 2 ? RING_CI * 272 + 16 : (FAM == 3 ? RING_CI * 144 + 16 : RING_CI * 176 + 32)));
    static constexpr int off_yq    = RING_ROWS * pitch_max;
    static constexpr int off_d     = off_yq + RING_CI * 8 * 32 * COLS_N;
    static constexpr int off_s     = off_d + RING_CI * COLS_N * 4;
    static constexpr int s_bytes   = FAM == 2 ? 0 : (FAM == 0 ? COLS_N * 16 : COLS_N * 32);         // per super-block: s32 (128 B), s16 / n32 (256 B), none
    static constexpr int stage     = off_s + RING_CI * s_bytes;                 // 63744 / 72448 / 86784 / 56064
    static constexpr int off_red   = stages * stage;                            // float red[RING_NWC][128]
    static constexpr int off_bar   = off_red + RING_NWC * 128 * 4;
    static constexpr int smem      = off_bar + 128;
};
static_assert(ring_geo<1>::smem <= 232448 && ring_geo<0>::smem <= 232448 && ring_geo<2>::smem <= 232448 && ring_geo<3>::smem <= 232448 && ring_geo<4>::smem <= 232448, "the rings must fit the 227 KB of shared memory a CTA can have");
// row pitch and bytes per 256 weights of a node of family FAM (only family 0 holds two formats)
// family 4 = Q6_K rows at 2-byte alignment: up to 14 bytes in front of a row's chunk, pitch 1712 = 48 mod 128 (eight distinct 16-byte bank groups)
// Q4_0 / Q8_0 fragments are 4-byte loads, 16 bytes per row group: pitch = 16 mod 128 spreads the 8 row groups over all banks
template <int FAM> __device__ __forceinline__ int ring_pitch(int type) {
    return FAM == 1 ? RING_CI * 210 : FAM == 4 ? RING_CI * 210 + 32 : (FAM == 2 ? RING_CI * 272 + 16 : (FAM == 3 ? RING_CI * 144 + 16 : (type == T_Q5_K ? RING_CI * 176 + 32 : RING_CI * 144 + 32)));
}
template <int FAM> __device__ __forceinline__ int ring_block_bytes(int type) {
    return FAM == 1 || FAM == 4 ? 210 : (FAM == 2 ? 272 : (FAM == 3 ? 144 : (type == T_Q5_K ? 176 : 144)));
}

__device__ __forceinline__ uint2 lds64(const uint8_t * p) { return *reinterpret_cast<const uint2 *>(p); }

template <int FAM>
__global__ void __launch_bounds__(RING_THREADS, 1) k_mmvq_cols_ring(const __grid_constant__ cols_args a) {
    using G = ring_geo<FAM>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t * full = reinterpret_cast<uint64_t *>(smem + G::off_bar), * empty = full + G::stages;
    uint64_t * red_full = empty + G::stages, * red_empty = red_full + 1;
    float (*red)[128] = reinterpret_cast<float (*)[128]>(smem + G::off_red);     // one buffer: the epilogue is done with a tile long before the next one ends
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < G::stages; s++) { mbar_init(&full[s], RING_NPROD); mbar_init(&empty[s], RING_NWC); }
        mbar_init(red_full, RING_NWC); mbar_init(red_empty, 1);
        mbar_fence_init();
    }
    __syncthreads();
    int ni = 0;
    if (warp >= RING_NWC && warp < RING_NWC + RING_NPROD) {
        // ---------------- producers: warp p copies rows 11 p .. 11 p + 10 of the tile's chunk (one per lane) and one activation plane
        const int pw = warp - RING_NWC, r_lo = 11 * pw, r = r_lo + lane;
        int stage = 0; uint32_t phase = 0;
        asm volatile("griddepcontrol.wait;" ::: "memory");        // the quantizer of this list has finished
        for (int u = blockIdx.x; u < a.units; u += gridDim.x) {
            while (ni + 1 < a.count && u >= a.nodes[ni + 1].unit0) ni++;
            const cols_node & nd = a.nodes[ni];
            const int row0 = (u - nd.unit0) * RING_ROWS, items = nd.k >> 8, bs = ring_block_bytes<FAM>(nd.type);
            const int rows = min(RING_ROWS, nd.m - row0), pitch = ring_pitch<FAM>(nd.type);
            const int my_rows = max(0, min(rows, r_lo + 11) - r_lo);
            const uint8_t * src = nd.w + (size_t) (row0 + r) * nd.row_stride;
            const cols_meta M = cols_meta_of(nd.ym, nd.k);
            for (int c0 = 0; c0 < items; c0 += RING_CI) {
                const uint32_t ci = (uint32_t) min(RING_CI, items - c0), bytes = ci * (uint32_t) bs;
                uint8_t * st = smem + stage * G::stage;
                mbar_wait(&empty[stage], phase ^ 1u);
                // third plane: sums of 32 (Q4_K / Q5_K: 128 B per super-block), -8 * sums (Q4_0: 256 B), sums of 16 (Q6_K: 256 B), none (Q8_0)
                const uint32_t abytes = ci * (pw == 0 ? 8 * 32 * COLS_N : (pw == 1 ? COLS_N * 4 : G::s_bytes));
                if (FAM == 4) {
                    // rows at 2-byte alignment (k / 256 not a multiple of 8): each row's chunk is copied as the 16-byte aligned range
                    // around it (the consumers add the same offset); the last bytes of the tensor's last row are not over-read:
                    // that one copy stops at the last whole 16 bytes and the lane stores the rest itself, before the arrive
                    uint32_t cp = 0, tot = 0;
                    const uint8_t * p0 = nullptr;
                    if (lane < 11 && r < rows) {
                        const uint8_t * p = src + (size_t) c0 * bs;
                        const uint32_t mis = (uint32_t) (reinterpret_cast<uintptr_t>(p) & 15);
                        p0 = p - mis; tot = mis + bytes;
                        const bool last = row0 + r == nd.m - 1 && c0 + RING_CI >= items;
                        cp = last ? (tot & ~15u) : ((tot + 15u) & ~15u);
                        for (uint32_t o = cp; o < tot; o += 2)
                            *reinterpret_cast<unsigned short *>(st + r * pitch + o) = __ldg(reinterpret_cast<const unsigned short *>(p0 + o));
                    }
                    const uint32_t total = __reduce_add_sync(0xFFFFFFFFu, cp);
                    __syncwarp();                                 // the lanes' tail stores are ordered before lane 0's (release) arrive
                    if (lane == 0) mbar_arrive_expect_tx(&full[stage], total + abytes);
                    __syncwarp();
                    if (cp) bulk_g2s(st + r * pitch, p0, cp, &full[stage]);
                } else {
                    if (lane == 0) mbar_arrive_expect_tx(&full[stage], bytes * (uint32_t) my_rows + abytes);
                    __syncwarp();
                    if (lane < 11 && r < rows) bulk_g2s(st + r * pitch, src + (size_t) c0 * bs, bytes, &full[stage]);
                }
                if (lane == 11) {
                    if (pw == 0)      bulk_g2s(st + G::off_yq, nd.yq + (size_t) c0 * (8 * 32 * COLS_N), abytes, &full[stage]);
                    else if (pw == 1) bulk_g2s(st + G::off_d, M.d + (size_t) c0 * COLS_N, abytes, &full[stage]);
                    else if (FAM == 1 || FAM == 4) bulk_g2s(st + G::off_s, M.s16 + (size_t) c0 * COLS_N * 4, abytes, &full[stage]);
                    else if (FAM == 2)          { }
                    else if (FAM == 3)          bulk_g2s(st + G::off_s, M.n32 + (size_t) c0 * 8 * COLS_N, abytes, &full[stage]);
                    else                        bulk_g2s(st + G::off_s, M.s32 + (size_t) c0 * COLS_N * 2, abytes, &full[stage]);
                }
                if (++stage == G::stages) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == RING_NWC + RING_NPROD) {
        // ---------------- epilogue: sum the eight partial tiles of each row half in warp order, store
        uint32_t rphase = 0;
        for (int u = blockIdx.x; u < a.units; u += gridDim.x, rphase ^= 1u) {
            while (ni + 1 < a.count && u >= a.nodes[ni + 1].unit0) ni++;
            const cols_node & nd = a.nodes[ni];
            const int row0 = (u - nd.unit0) * RING_ROWS;
            mbar_wait(red_full, rphase);
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const int e = lane + 32 * j, half = e >> 7, idx = e & 127, l = idx >> 2, i = idx & 3;
                float v = 0.0f;
#pragma unroll
                for (int w = 0; w < 8; w++) v += red[half * 8 + w][idx];
                // c fragment layout: element i of lane l = (g, t): row g + 8 (i >> 1), column 2t + (i & 1)
                const int row = row0 + 16 * half + (l >> 2) + 8 * (i >> 1), col = 2 * (l & 3) + (i & 1);
                if (row < nd.m && col < nd.n) nd.dst[(size_t) col * nd.m + row] = v;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive_cta(red_empty);
        }
    } else {
        // ---------------- consumers
        const int g = lane >> 2, t = lane & 3, half = warp >> 3, slot = warp & 7;
        int stage = 0; uint32_t phase = 0, rphase = 0;
        for (int u = blockIdx.x; u < a.units; u += gridDim.x, rphase ^= 1u) {
            while (ni + 1 < a.count && u >= a.nodes[ni + 1].unit0) ni++;
            const cols_node & nd = a.nodes[ni];
            const int items = nd.k >> 8;
            const bool q5 = nd.type == T_Q5_K;
            const int bs = ring_block_bytes<FAM>(nd.type), pitch = ring_pitch<FAM>(nd.type);
            float acc[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
            for (int c0 = 0; c0 < items; c0 += RING_CI) {
                mbar_wait(&full[stage], phase);
                if (c0 + slot < items) {
                    const uint8_t * st = smem + stage * G::stage;
                    const uint8_t * ra = st + (16 * half + g) * pitch + slot * bs, * rb = ra + 8 * pitch;
                    const uint2 dyb = lds64(st + G::off_d + slot * (COLS_N * 4) + 8 * t);                   // columns 2t, 2t + 1
                    const float2 dy = make_float2(__uint_as_float(dyb.x), __uint_as_float(dyb.y));
                    // (columns >= n hold whatever the workspace held: their results are never stored)
                    if (FAM == 2 || FAM == 3) {
                        // Q4_0 / Q8_0: eight 32-weight blocks with an f16 scale each; lane (g, t) takes quant bytes 4t..4t+3 (k slots
                        // 4t..) and, Q4_0: their high nibbles / Q8_0: bytes 16 + 4t.. (k slots 16 + 4t..) -- the activation
                        // bytes of the same positions.  Blocks of even index sit 2 bytes off word alignment: funnel shifts.
                        constexpr bool Q8 = FAM == 2;
                        constexpr int BB = Q8 ? 34 : 18;
                        // (the list's quantizer stored them permuted -- cols_node::perm -- so that the two words are one 8-byte load)
                        const uint8_t * yqs = st + G::off_yq + (slot * 8) * (32 * COLS_N) + g * 32 + 8 * t;
                        float sbs[4] = { 0.0f, 0.0f, 0.0f, 0.0f };          // the 256 weights' sums before the activations' scale (one per super-block)
                        auto blk = [&](auto Bc) {
                            constexpr int b = decltype(Bc)::value;
                            constexpr int o = BB * b + 2;                   // the block's quants; (o & 3) == 2 for even b
                            uint32_t wa0, wb0, wa1 = 0, wb1 = 0;
                            if (b & 1) {
                                wa0 = lds32(ra + o + 4 * t); wb0 = lds32(rb + o + 4 * t);
                                if (Q8) { wa1 = lds32(ra + o + 16 + 4 * t); wb1 = lds32(rb + o + 16 + 4 * t); }
                            } else {
                                wa0 = __funnelshift_r(lds32(ra + o - 2 + 4 * t), lds32(ra + o + 2 + 4 * t), 16);
                                wb0 = __funnelshift_r(lds32(rb + o - 2 + 4 * t), lds32(rb + o + 2 + 4 * t), 16);
                                if (Q8) {
                                    wa1 = __funnelshift_r(lds32(ra + o + 14 + 4 * t), lds32(ra + o + 18 + 4 * t), 16);
                                    wb1 = __funnelshift_r(lds32(rb + o + 14 + 4 * t), lds32(rb + o + 18 + 4 * t), 16);
                                }
                            }
                            const float da = f16_bits_to_f32(lds16(ra + BB * b)), db = f16_bits_to_f32(lds16(rb + BB * b));
                            const uint2 bb = lds64(yqs + b * (32 * COLS_N));
                            const uint32_t b0 = bb.x, b1 = bb.y;
                            int c[4];
                            if (Q8) mma_s8s8_16x8x32(c, wa0, wb0, wa1, wb1, b0, b1, COLS_I2F_BIAS);
                            else {
                                const uint2 nn = lds64(st + G::off_s + ((slot * 8 + b) * COLS_N + 2 * t) * 4);      // -8 * sum32, columns 2t, 2t + 1
                                mma_u8s8_16x8x32_c(c, wa0 & 0x0F0F0F0Fu, wb0 & 0x0F0F0F0Fu, (wa0 >> 4) & 0x0F0F0F0Fu, (wb0 >> 4) & 0x0F0F0F0Fu,
                                                   b0, b1, (int) nn.x, (int) nn.y);
                            }
                            // c = bias + the block's integer dot: its bits are the float 1.5 * 2^23 + dot, one exact FADD away from (float) dot
                            sbs[0] = fmaf(da, __int_as_float(c[0]) - COLS_I2F_BIASF, sbs[0]); sbs[1] = fmaf(da, __int_as_float(c[1]) - COLS_I2F_BIASF, sbs[1]);
                            sbs[2] = fmaf(db, __int_as_float(c[2]) - COLS_I2F_BIASF, sbs[2]); sbs[3] = fmaf(db, __int_as_float(c[3]) - COLS_I2F_BIASF, sbs[3]);
                        };
                        blk(std::integral_constant<int, 0>()); blk(std::integral_constant<int, 1>()); blk(std::integral_constant<int, 2>());
                        blk(std::integral_constant<int, 3>()); blk(std::integral_constant<int, 4>()); blk(std::integral_constant<int, 5>());
                        blk(std::integral_constant<int, 6>()); blk(std::integral_constant<int, 7>());
                        acc[0] = fmaf(dy.x, sbs[0], acc[0]); acc[1] = fmaf(dy.y, sbs[1], acc[1]);
                        acc[2] = fmaf(dy.x, sbs[2], acc[2]); acc[3] = fmaf(dy.y, sbs[3], acc[3]);
                    } else if (FAM == 1 || FAM == 4) {
                        const uint2 s16 = lds64(st + G::off_s + (slot * COLS_N + g) * 32 + 8 * t);          // column g, groups 4t..4t+3
                        const int8_t * yq = reinterpret_cast<const int8_t *>(st + G::off_yq) + g * 32;
                        cols_item6 ia, ib;
                        if (FAM == 4) {
                            // the row's chunk sits `mis` bytes into its stage row (the producer copied the aligned range around it);
                            // rows g and g + 8 are a multiple of 16 bytes apart
                            const int row0 = (u - nd.unit0) * RING_ROWS;
                            const uint32_t mis = (uint32_t) (reinterpret_cast<uintptr_t>(nd.w + (size_t) (row0 + 16 * half + g) * nd.row_stride + (size_t) c0 * bs) & 15);
                            cols_load6_var(ia, ra + mis, t); cols_load6_var(ib, rb + mis, t);
                        }
                        else if (slot & 1) { cols_load6<true, true>(ia, ra, t);  cols_load6<true, true>(ib, rb, t); }
                        else               { cols_load6<false, true>(ia, ra, t); cols_load6<false, true>(ib, rb, t); }
                        cols_item6_mma<true, true>(ia, ib, slot, t, yq, dy, s16, true, acc);
                    } else {
                        uint2 b[8];
#pragma unroll
                        for (int q = 0; q < 8; q++) b[q] = lds64(st + G::off_yq + (slot * 8 + q) * (32 * COLS_N) + g * 32 + 8 * t);
                        uint2 s32 = make_uint2(0u, 0u);
                        if (t < 2) s32 = lds64(st + G::off_s + (slot * COLS_N + g) * 16 + 8 * t);           // column g, sub-blocks 4t..4t+3
                        cols_item ia, ib;
                        ia.hdr = lds128(ra); ib.hdr = lds128(rb);
                        const int qo = (q5 ? 48 : 16) + 8 * t;
#pragma unroll
                        for (int p = 0; p < 4; p++) { ia.q[p] = lds64(ra + qo + 32 * p); ib.q[p] = lds64(rb + qo + 32 * p); }
                        if (q5) {
                            ia.h = lds64(ra + 16 + 8 * t); ib.h = lds64(rb + 16 + 8 * t);
                            cols_sb_mma<true>(ia, ib, b, dy, s32, t, acc, nullptr, nullptr, false);
                        } else {
                            cols_sb_mma<false>(ia, ib, b, dy, s32, t, acc, nullptr, nullptr, false);
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive_cta(&empty[stage]);
                if (++stage == G::stages) { stage = 0; phase ^= 1u; }
            }
            mbar_wait(red_empty, rphase ^ 1u);                   // the epilogue has read the previous tile
            *reinterpret_cast<float4 *>(&red[warp][lane * 4]) = make_float4(acc[0], acc[1], acc[2], acc[3]);
            __syncwarp();
            if (lane == 0) mbar_arrive_cta(red_full);
        }
    }
}

// ---------------------------------------------------------------- host side
static inline bool cols_node_ok(int type, const uint8_t * w, int64_t row_stride, int64_t m, int64_t k, const float * x,
                                int64_t x_stride_bytes, int64_t n) {
    static const int ring_on = env_int("GGML_B200_COLS_RING", 1);
    if (type != T_Q4_K && type != T_Q5_K && type != T_Q6_K && !((type == T_Q4_0 || type == T_Q8_0) && ring_on)) return false;
    if (n < 2 || n > COLS_N || m < 1 || k < 256 || k % 256 != 0) return false;
    if (m >= (int64_t) 1 << 31 || k >= (int64_t) 1 << 24) return false;
    if ((reinterpret_cast<uintptr_t>(w) & 15) || (reinterpret_cast<uintptr_t>(x) & 3) || (x_stride_bytes & 3)) return false;
    // rows that are multiples of 16 bytes; Q6_K (210-byte super-blocks: only when k / 256 is a multiple of 8) also at 2-byte alignment,
    // where the ring copies the aligned range around each row's chunk (cols_node::unal) -- GGML_B200_COLS_UNAL=0: as one entry per column
    static const int unal_on = env_int("GGML_B200_COLS_UNAL", 1);
    if (type == T_Q6_K ? ((row_stride & 1) || (!(unal_on && ring_on) && ((row_stride & 15) || (k / 256) % 8 != 0))) : (row_stride & 15) != 0) return false;
    return row_stride >= k / 256 * cols_block_bytes(type);
}
static inline size_t cols_ws_bytes(int64_t k, int64_t n) {
    (void) n;
    // int8 [k/32][8][32]  +  per super-block: f32 d [8], sums of 32 (2 x 8 bytes) [8], sums of 16 (4 x 8 bytes) [8] (cols_meta)
    // + -8 * sums of 32 as int32 [8 sub-blocks][8]
    return (((size_t) COLS_N * k + 127) & ~(size_t) 127) + (((size_t) (k / 256) * COLS_N * (4 + 16 + 32 + 32) + 127) & ~(size_t) 127);
}

struct cols_builder {
    cols_args args;       // tuning (GGML_B200_COLS_RING=0): k_mmvq_cols, 16-row tiles
    cols_args ring;       // Q4_K / Q5_K nodes: k_mmvq_cols_ring<0>, 32-row tiles
    cols_args ring6;      // Q6_K nodes: k_mmvq_cols_ring<1>
    cols_args ring8;      // Q8_0 nodes: k_mmvq_cols_ring<2>
    cols_args ring4;      // Q4_0 nodes: k_mmvq_cols_ring<3>
    cols_args ring6u;     // Q6_K nodes whose rows are not 16-byte multiples: k_mmvq_cols_ring<4>
    // the activations quantized last (nodes of either list that read the same x share them: q/k/v, gate/up)
    const float * last_x; int64_t last_k, last_n, last_stride; int8_t * last_yq; uint8_t * last_ym; int last_perm;
    cols_builder() { reset(); }
    void reset() {
        args.count = 0; args.units = 0; args.kmax = 0; args.nq = 0;
        ring.count = 0; ring.units = 0; ring.kmax = 0; ring.nq = 0;
        ring6.count = 0; ring6.units = 0; ring6.kmax = 0; ring6.nq = 0;
        ring8.count = 0; ring8.units = 0; ring8.kmax = 0; ring8.nq = 0;
        ring4.count = 0; ring4.units = 0; ring4.kmax = 0; ring4.nq = 0;
        ring6u.count = 0; ring6u.units = 0; ring6u.kmax = 0; ring6u.nq = 0;
        last_x = nullptr; last_k = last_n = last_stride = 0; last_yq = nullptr; last_ym = nullptr; last_perm = 0;
    }
    bool full() const { return args.count >= COLS_MAX_NODES || ring.count >= COLS_MAX_NODES || ring6.count >= COLS_MAX_NODES || ring8.count >= COLS_MAX_NODES || ring4.count >= COLS_MAX_NODES || ring6u.count >= COLS_MAX_NODES; }
    bool empty() const { return args.count == 0 && ring.count == 0 && ring6.count == 0 && ring8.count == 0 && ring4.count == 0 && ring6u.count == 0; }
    int  count() const { return args.count + ring.count + ring6.count + ring8.count + ring4.count + ring6u.count; }
    bool add(int type, const uint8_t * w, int64_t row_stride, int64_t m, int64_t k, const float * x, int64_t x_stride_bytes,
             int64_t n, float * dst, ws_cursor & cur) {
        static const int use_ring = env_int("GGML_B200_COLS_RING", 1);
        const bool unal = type == T_Q6_K && ((row_stride & 15) || (k / 256) % 8 != 0);
        cols_args & A = !use_ring ? args : (type == T_Q6_K ? (unal ? ring6u : ring6) : (type == T_Q8_0 ? ring8 : (type == T_Q4_0 ? ring4 : ring)));
        const int tile = use_ring ? RING_ROWS : 16;
        cols_node & nd = A.nodes[A.count];
        const int perm = use_ring && (type == T_Q4_0 || type == T_Q8_0 || type == T_Q6_K);
        if (last_x == x && last_k == k && last_n == n && last_stride == x_stride_bytes && last_perm == perm) {
            nd.yq = last_yq; nd.ym = last_ym; nd.quantize = 0;     // same activations as the node before (gate/up, q/k/v)
        } else {
            const size_t need = cols_ws_bytes(k, n);
            if (need > cur.left) return false;
            nd.yq = reinterpret_cast<int8_t *>(cur.p);
            nd.ym = cur.p + (((size_t) COLS_N * k + 127) & ~(size_t) 127);
            nd.quantize = 1; cur.p += need; cur.left -= need;
            A.qlist[A.nq++] = (unsigned short) A.count;
            last_x = x; last_k = k; last_n = n; last_stride = x_stride_bytes; last_yq = nd.yq; last_ym = nd.ym; last_perm = perm;
        }
        nd.w = w; nd.x = x; nd.dst = dst; nd.row_stride = row_stride; nd.m = (int) m; nd.k = (int) k; nd.n = (int) n;
        nd.x_stride = (int) (x_stride_bytes / 4); nd.unit0 = A.units; nd.type = (unsigned char) type; nd.perm = (unsigned char) perm;
        nd.unal = (unsigned char) unal;
        A.units += (int) ((m + tile - 1) / tile);
        if (nd.k > A.kmax) A.kmax = nd.k;
        A.count++;
        return true;
    }
};

static int cols_launch(cols_builder & cb, cudaStream_t st) {
    if (cb.empty()) return B200_OK;
    const dev_caps & caps = current_caps();
    static const int pdl = env_int("GGML_B200_FUSED_PDL", 1);
    // 1. the activations of both lists (a node that shares them with an earlier node of either list does nothing)
    for (cols_args * A : { &cb.ring, &cb.ring6, &cb.ring8, &cb.ring4, &cb.ring6u, &cb.args }) {
        if (A->nq == 0) continue;
        const dim3 grid((unsigned) ((A->kmax / 256 + 7) / 8), (unsigned) A->nq, COLS_N);
        k_quantize_cols<<<grid, 256, 0, st>>>(*A);
        b200_count_launch(1);
    }
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    // 2. the tiles through the TMA ring, one launch per family (starts while the quantizers drain)
    auto ring_launch = [&](auto kern, size_t smem, cols_args & A, smem_limit & lim) -> int {
        if (const int rc = lim.ensure(kern, smem)) return rc;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned) min(caps.sm_count, A.units)); cfg.blockDim = dim3(RING_THREADS);
        cfg.dynamicSmemBytes = smem; cfg.stream = st;
        cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
        const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, A);
        if (e != cudaSuccess) return -(int) e;
        b200_count_launch(1);
        return B200_OK;
    };
    static smem_limit lim0, lim1, lim2, lim3, lim4;
    if (cb.ring.count)  { if (const int rc = ring_launch(k_mmvq_cols_ring<0>, ring_geo<0>::smem, cb.ring, lim0)) return rc; }
    if (cb.ring6.count) { if (const int rc = ring_launch(k_mmvq_cols_ring<1>, ring_geo<1>::smem, cb.ring6, lim1)) return rc; }
    if (cb.ring8.count) { if (const int rc = ring_launch(k_mmvq_cols_ring<2>, ring_geo<2>::smem, cb.ring8, lim2)) return rc; }
    if (cb.ring4.count) { if (const int rc = ring_launch(k_mmvq_cols_ring<3>, ring_geo<3>::smem, cb.ring4, lim3)) return rc; }
    if (cb.ring6u.count) { if (const int rc = ring_launch(k_mmvq_cols_ring<4>, ring_geo<4>::smem, cb.ring6u, lim4)) return rc; }
    // 3. tuning fallback: per-lane loads behind L2 prefetches
    if (cb.args.count) {
        cols_args & a = cb.args;
        int ctas = caps.sm_count * COLS_CTAS;
        if (ctas > a.units) ctas = a.units;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned) ctas); cfg.blockDim = dim3(COLS_NW * 32);
        cfg.dynamicSmemBytes = 0; cfg.stream = st;
        cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
        static const int ks1 = env_int("GGML_B200_COLS_KS", 8) == 1;      // tuning: whole tiles per warp instead of the K split
        const cudaError_t e = ks1 ? cudaLaunchKernelEx(&cfg, k_mmvq_cols<1>, a) : cudaLaunchKernelEx(&cfg, k_mmvq_cols<8>, a);
        if (e != cudaSuccess) return -(int) e;
        b200_count_launch(1);
    }
    return B200_OK;
}

} // namespace b200
