// mmvq_cols.cuh — decode at batch 2..8 (MUL_MAT with a few activation columns), Q4_K / Q5_K / Q6_K weights:
// ONE launch pair per list of independent nodes, int8 tensor-core MMA on the weight nibbles as they are.
//
// The reference sends n <= 8 through mul_mat_vec_q with ncols_dst template instances (src/ggml-cuda/mmvq.cu:139-226,
// one dp4a chain per column) and larger n to the MMQ tiles (mmq.cuh).  On B200 the dp4a chains make the decode
// list kernel issue-bound as soon as there is more than one column (measured: time = n x the one-column time even
// with the weights in L2), so here a warp multiplies a 16-row x 32-weight sub-block by the 8 (padded) columns
// with one mma.sync.m16n8k32.s8 and applies the block scales on the 16x8 int32 result:
//   * the k4 nibble layout IS the A fragment: lane (g, t) takes bytes 8t..8t+7 of a 32-byte quant chunk of rows g
//     and g + 8 (the k order inside a 32-block is free as long as A and B agree), whose low / high nibbles are the
//     sub-blocks 2j / 2j+1;
//   * activations are quantized once per list to int8, the 8 columns of a 32-block side by side, + (d, d * sum)
//     pairs per 32-block (q8_1 arithmetic), and read as B fragments from global memory (L1-resident: 10 k bytes);
//   * work unit = (node, 16-row tile), dealt round-robin to two CTAs per SM; the 8 warps of a CTA split K in
//     256-weight super-blocks; whole tiles are requested into L2 two units ahead by bulk prefetches, the warps'
//     own loads then hit L2, and the registers hold the activation fragments half a super-block ahead of their
//     MMAs; partial 16x8 tiles meet in shared memory and are added in fixed warp order: a row's value does not
//     depend on the grid.
// Measured (profiles/r01o_bs_probe.txt, r01p_ncu_k_mmvq_cols_summary.txt): 2.85 TB/s on the weight bytes of the
// Llama-2-7B Q4_K stack at 8 columns (44 % of the measured HBM peak, 6100 tok/s = 4x one-column decode); issue-bound
// (441 instructions per 16-row super-block column, 61 % issue-active with 4 warps per scheduler).
#pragma once
#include "mmvq_batch.cuh"

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
    float2        * yds;      // workspace: (d, d * sum) [k / 32][8]
    int64_t row_stride;
    int   m, k, n, x_stride, unit0;     // unit0: index of this node's first 16-row tile in the launch
    unsigned char type, quantize, pad[2];
};
static_assert(sizeof(cols_node) == 72, "cols_node is part of the kernel parameter block");
static_assert(COLS_NW == 8, "k_mmvq_cols<8> splits K over the 8 warps of a CTA");

struct cols_args {
    int count, units, kmax, pad;
    cols_node nodes[COLS_MAX_NODES];
};

// third activation plane, behind yds: int32 [k / 16][8] = -32 * (sum of the 16 int8 activations)
__device__ __host__ __forceinline__ int * cols_yns(const cols_node & nd) { return reinterpret_cast<int *>(nd.yds + (size_t) (nd.k / 32) * COLS_N); }

// ---------------------------------------------------------------- activations of the list: f32 -> int8 columns
// grid (ceil(kmax / 32 / 8), count, 8), block 256: one warp per 32 values of one column (q8_1 arithmetic as quantize.cu)
__global__ void __launch_bounds__(256) k_quantize_cols(const __grid_constant__ cols_args a) {
    const cols_node & nd = a.nodes[blockIdx.y];
    const int kb = blockIdx.x * 8 + (threadIdx.x >> 5), c = blockIdx.z;
    const int lane = threadIdx.x & 31;
    if (!nd.quantize || kb * 32 >= nd.k) return;
    if (c >= nd.n) {                                      // padding columns: zero scale pairs, q never read as non-zero
        if (lane == 0) nd.yds[kb * COLS_N + c] = make_float2(0.0f, 0.0f);
        if (lane < 2) cols_yns(nd)[(2 * kb + lane) * COLS_N + c] = 0;
        return;
    }
    const float x = __ldg(nd.x + (size_t) c * nd.x_stride + kb * 32 + lane);
    const float amax = warp_max(fabsf(x));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const float v  = roundf(__fmul_rn(x, id));
    const int   qi = fabsf(v) < 2147483648.0f ? (int) v : (int) 0x80000000;
    const int   q8 = (int) (int8_t) (qi & 0xFF);
    nd.yq[((size_t) kb * COLS_N + c) * 32 + lane] = (int8_t) q8;
    int sum = q8;
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, off);
    // Q6_K weights are (q - 32) * scale per 16 weights: -32 * (sum of the 16 activations) is the C input of that MMA
    if ((lane & 15) == 0) cols_yns(nd)[(2 * kb + (lane >> 4)) * COLS_N + c] = -32 * sum;
    sum += __shfl_xor_sync(0xFFFFFFFFu, sum, 16);
    if (lane == 0) {
        const float dh = __half2float(__float2half_rn(d));
        nd.yds[kb * COLS_N + c] = make_float2(dh, dh * (float) sum);
    }
}

// ---------------------------------------------------------------- the MMA
__device__ __forceinline__ void mma_s8_16x8x32(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
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

__device__ __forceinline__ int cols_block_bytes(int type) { return type == T_Q6_K ? 210 : (type == T_Q5_K ? 176 : 144); }

// HBM -> L2 prefetch of a whole weight row (16-byte aligned, a multiple of 16 bytes): no registers, no completion
__device__ __forceinline__ void l2_prefetch_bulk(const void * p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p), "r"(bytes) : "memory");
}

// The words of one 128-weight item (half a super-block) of one row that lane (g, t) needs.  The MMA's k order
// inside a 32-block is free as long as A and B agree: lane t takes bytes 8t..8t+7 of the 32-byte quant chunk
// (one 8-byte load) as its k slots 4t..4t+3 and 16+4t..16+4t+3, and the same bytes of the activation block.
struct cols_item {            // one 256-weight super-block of one row, as lane (g, t) needs it
    uint4 hdr;                // d, dmin (f16 x 2) + 12 scale bytes
    uint2 q[4];               // [32-byte quant chunk] bytes 8t..8t+7
    uint2 h;                  // Q5_K: high-bit bytes 8t..8t+7
};

__device__ __forceinline__ void cols_load(cols_item & it, const uint8_t * row, int item, int t, bool q5) {
    const uint8_t * blk = row + (size_t) item * (q5 ? 176 : 144);
    it.hdr = ldg_stream128(blk);
    const uint8_t * qs = blk + (q5 ? 48 : 16) + 8 * t;
#pragma unroll
    for (int p = 0; p < 4; p++) it.q[p] = ldg_stream64(qs + 32 * p);
    if (q5) it.h = ldg_stream64(blk + 16 + 8 * t);
}

// d * sc and dmin * m of the item's four sub-blocks (get_scale_min_k4 on four bytes at a time)
__device__ __forceinline__ void cols_scales(const uint4 & hdr, int half, float (&dsc)[4], float (&dmn)[4]) {
    const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&hdr.x));
    uint32_t sc, mn;
    if (half == 0) { sc = hdr.y & 0x3F3F3F3Fu; mn = hdr.z & 0x3F3F3F3Fu; }
    else {
        sc = (hdr.w & 0x0F0F0F0Fu) | ((hdr.y >> 2) & 0x30303030u);
        mn = ((hdr.w >> 4) & 0x0F0F0F0Fu) | ((hdr.z >> 2) & 0x30303030u);
    }
#pragma unroll
    for (int q = 0; q < 4; q++) {
        dsc[q] = dm.x * (float) ((sc >> (8 * q)) & 0xFFu);
        dmn[q] = dm.y * (float) ((mn >> (8 * q)) & 0xFFu);
    }
}

// B fragments and (d, d * sum) pairs of four consecutive 32-blocks (half a super-block), requested together
struct cols_acts { uint2 b[4]; float4 ds[4]; };
__device__ __forceinline__ void cols_load_acts(cols_acts & y, const int8_t * yq_col, const float2 * yds, int blk0, int t, bool col_ok) {
    const int8_t * yq = yq_col + (size_t) blk0 * (32 * COLS_N) + 8 * t;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        y.b[q] = make_uint2(0u, 0u);
        if (col_ok) y.b[q] = __ldg(reinterpret_cast<const uint2 *>(yq + q * (32 * COLS_N)));
        y.ds[q] = __ldg(reinterpret_cast<const float4 *>(yds + (size_t) (blk0 + q) * COLS_N + 2 * t));   // columns 2t, 2t + 1
    }
}

// Half a super-block of rows g and g + 8 against the 8 columns: 4 MMAs (one per 32-weight sub-block), block
// scales applied to the 16x8 int32 result.
__device__ __forceinline__ void cols_half_mma(const cols_item & ia, const cols_item & ib, const int half, bool q5,
                                              const cols_acts & y, float (&acc)[4]) {
    float dscA[4], dmnA[4], dscB[4], dmnB[4];
    cols_scales(ia.hdr, half, dscA, dmnA);
    cols_scales(ib.hdr, half, dscB, dmnB);
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const int p = half * 2 + (q >> 1), sh = (q & 1) * 4;
        uint32_t a0 = (ia.q[p].x >> sh) & 0x0F0F0F0Fu, a1 = (ib.q[p].x >> sh) & 0x0F0F0F0Fu;
        uint32_t a2 = (ia.q[p].y >> sh) & 0x0F0F0F0Fu, a3 = (ib.q[p].y >> sh) & 0x0F0F0F0Fu;
        if (q5) {
            const int s = half * 4 + q;                                  // bit s of qh belongs to sub-block s
            a0 |= ((ia.h.x >> s) & 0x01010101u) << 4; a1 |= ((ib.h.x >> s) & 0x01010101u) << 4;
            a2 |= ((ia.h.y >> s) & 0x01010101u) << 4; a3 |= ((ib.h.y >> s) & 0x01010101u) << 4;
        }
        int c[4] = { 0, 0, 0, 0 };
        mma_s8_16x8x32(c, a0, a1, a2, a3, y.b[q].x, y.b[q].y);
        const float4 ds = y.ds[q];
        acc[0] += dscA[q] * (ds.x * (float) c[0]) - dmnA[q] * ds.y;
        acc[1] += dscA[q] * (ds.z * (float) c[1]) - dmnA[q] * ds.w;
        acc[2] += dscB[q] * (ds.x * (float) c[2]) - dmnB[q] * ds.y;
        acc[3] += dscB[q] * (ds.z * (float) c[3]) - dmnB[q] * ds.w;
    }
}

// ---------------------------------------------------------------- Q6_K: one m16n8k16 MMA per 16-weight scale group
// block = ql[128] qh[64] sc[16] (int8) d (f16) = 210 bytes: blocks sit at 2-byte alignment.  With w and the row
// stride multiples of 4 the blocks of even index are 4-byte aligned and those of odd index are 2 bytes off, the
// same for every lane: odd blocks read the two aligned words around each word and funnel-shift.
// Lane (g, t) takes bytes 4t..4t+3 of each 16-byte piece: for half h, piece s, word j of ql serves the groups
// r = j (low nibbles) and r = j + 2 (high nibbles); the qh word serves r = 0..3 (2 bits each).
__device__ __forceinline__ void mma_s8_16x8x16(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%0, %1, %2, %3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
}
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
template <bool ODD>
__device__ __forceinline__ uint32_t cols_ld6(const uint8_t * blk, int off) {
    if (!ODD) return ldg_stream32(blk + off);
    return __funnelshift_r(ldg_stream32(blk + off - 2), ldg_stream32(blk + off + 2), 16);
}
template <bool ODD>
__device__ __forceinline__ void cols_load6(cols_item6 & it, const uint8_t * blk, int t) {
#pragma unroll
    for (int h = 0; h < 2; h++)
#pragma unroll
        for (int s = 0; s < 2; s++) {
            it.ql[h][0][s] = cols_ld6<ODD>(blk, 64 * h + 16 * s + 4 * t);
            it.ql[h][1][s] = cols_ld6<ODD>(blk, 64 * h + 32 + 16 * s + 4 * t);
            it.qh[h][s]    = cols_ld6<ODD>(blk, 128 + 32 * h + 16 * s + 4 * t);
        }
    if (!ODD) {
#pragma unroll
        for (int i = 0; i < 4; i++) it.sc[i] = ldg_stream32(blk + 192 + 4 * i);
    } else {
        uint32_t w[5];
#pragma unroll
        for (int i = 0; i < 5; i++) w[i] = ldg_stream32(blk + 190 + 4 * i);
#pragma unroll
        for (int i = 0; i < 4; i++) it.sc[i] = __funnelshift_r(w[i], w[i + 1], 16);
    }
    it.d = *reinterpret_cast<const unsigned short *>(blk + 208);
}
// sign-extended byte i of w as float
__device__ __forceinline__ float cols_s8f(uint32_t w, int i) {
    return (float) ((int) (w << (24 - 8 * i)) >> 24);      // (__byte_perm ignores prmt's sign-replicate selector bit)
}
__device__ __forceinline__ void cols_item6_mma(const cols_item6 & ia, const cols_item6 & ib, int item, int g, int t,
                                               const int8_t * yq_col, const float2 * yds, const int * yns, bool col_ok, float (&acc)[4]) {
    const float dA = f16_bits_to_f32(ia.d), dB = f16_bits_to_f32(ib.d);
#pragma unroll
    for (int h = 0; h < 2; h++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const int B = item * 8 + 4 * h + r;                         // the 32-block of the activations
            const float4 ds = __ldg(reinterpret_cast<const float4 *>(yds + (size_t) B * COLS_N + 2 * t));     // .x, .z: d of columns 2t, 2t + 1
            float tA0 = 0.0f, tA1 = 0.0f, tB0 = 0.0f, tB1 = 0.0f;
#pragma unroll
            for (int s = 0; s < 2; s++) {
                const uint32_t a0 = ((ia.ql[h][r & 1][s] >> (4 * (r >> 1))) & 0x0F0F0F0Fu) | (((ia.qh[h][s] >> (2 * r)) & 0x03030303u) << 4);
                const uint32_t a1 = ((ib.ql[h][r & 1][s] >> (4 * (r >> 1))) & 0x0F0F0F0Fu) | (((ib.qh[h][s] >> (2 * r)) & 0x03030303u) << 4);
                uint32_t b0 = 0u;
                if (col_ok) b0 = __ldg(reinterpret_cast<const uint32_t *>(yq_col + (size_t) B * (32 * COLS_N) + 16 * s + 4 * t));
                const int2 ns = __ldg(reinterpret_cast<const int2 *>(yns + (size_t) (2 * B + s) * COLS_N + 2 * t));
                int c[4] = { ns.x, ns.y, ns.x, ns.y };                  // -32 * sum16: the weights enter as q, not q - 32
                mma_s8_16x8x16(c, a0, a1, b0);
                const int si = 8 * h + 2 * r + s;                       // scale of this 16-weight group
                const float fA = cols_s8f(ia.sc[si >> 2], si & 3), fB = cols_s8f(ib.sc[si >> 2], si & 3);
                tA0 += fA * (float) c[0]; tA1 += fA * (float) c[1];
                tB0 += fB * (float) c[2]; tB1 += fB * (float) c[3];
            }
            acc[0] += (dA * ds.x) * tA0; acc[1] += (dA * ds.z) * tA1;
            acc[2] += (dB * ds.x) * tB0; acc[3] += (dB * ds.z) * tB1;
        }
    }
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
    for (; u < a.units; u += ustride, par ^= 1) {
        while (ni + 1 < a.count && u >= a.nodes[ni + 1].unit0) ni++;
        const cols_node & nd = a.nodes[ni];
        const int row0 = (u - nd.unit0) * 16, items = nd.k >> 8;
        const bool q5 = nd.type == T_Q5_K;
        const int ra = min(row0 + g, nd.m - 1), rb = min(row0 + g + 8, nd.m - 1);      // tail rows: recomputed, never stored
        const uint8_t * pa = nd.w + (size_t) ra * nd.row_stride;
        const uint8_t * pb = nd.w + (size_t) rb * nd.row_stride;
        if (KS > 1) prefetch_unit(u + PD * ustride, nip);
        else if (u + ustride < a.units) { while (nip + 1 < a.count && u + ustride >= a.nodes[nip + 1].unit0) nip++; }   // nip: the next tile's node
        const int8_t * yq_col = nd.yq + g * 32;
        const float2 * yds = nd.yds;
        const bool col_ok = g < nd.n;
        float acc[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
        if (nd.type == T_Q6_K) {
            const int * yns = cols_yns(nd);
            for (int it = kw; it < items; it += KS) {
                cols_item6 ia, ib;
                if (it & 1) { cols_load6<true>(ia, pa + (size_t) it * 210, t);  cols_load6<true>(ib, pb + (size_t) it * 210, t); }
                else        { cols_load6<false>(ia, pa + (size_t) it * 210, t); cols_load6<false>(ib, pb + (size_t) it * 210, t); }
                cols_item6_mma(ia, ib, it, g, t, yq_col, yds, yns, col_ok, acc);
            }
        } else
        for (int it = kw; it < items; it += KS) {
            if (KS == 1 && (it & 7) == 0) {
                if (it + 8 < items) prefetch_chunk(nd, row0, it + 8);
                else if (u + ustride < a.units) prefetch_chunk(a.nodes[nip], (u + ustride - a.nodes[nip].unit0) * 16, 0);
            }
            cols_item ia, ib;
            cols_acts y0, y1;
            cols_load(ia, pa, it, t, q5); cols_load(ib, pb, it, t, q5);
            cols_load_acts(y0, yq_col, yds, it * 8, t, col_ok);
            cols_load_acts(y1, yq_col, yds, it * 8 + 4, t, col_ok);
            cols_half_mma(ia, ib, 0, q5, y0, acc);
            cols_half_mma(ia, ib, 1, q5, y1, acc);
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

// ---------------------------------------------------------------- host side
static inline bool cols_node_ok(int type, const uint8_t * w, int64_t row_stride, int64_t m, int64_t k, const float * x,
                                int64_t x_stride_bytes, int64_t n) {
    if (type != T_Q4_K && type != T_Q5_K && type != T_Q6_K) return false;
    if (n < 2 || n > COLS_N || m < 1 || k < 256 || k % 256 != 0) return false;
    if (m >= (int64_t) 1 << 31 || k >= (int64_t) 1 << 24) return false;
    if ((reinterpret_cast<uintptr_t>(w) & 15) || (row_stride & 15) || (reinterpret_cast<uintptr_t>(x) & 3) || (x_stride_bytes & 3)) return false;
    return row_stride >= k / 256 * (type == T_Q6_K ? 210 : (type == T_Q5_K ? 176 : 144));
}
static inline size_t cols_ws_bytes(int64_t k, int64_t n) {
    (void) n;
    // int8 [k/32][8][32]  +  (d, d * sum) [k/32][8]  +  -32 * sum16 [k/16][8]
    return (((size_t) COLS_N * k + 127) & ~(size_t) 127) + (((size_t) (k / 32) * COLS_N * (sizeof(float2) + 2 * sizeof(int)) + 127) & ~(size_t) 127);
}

struct cols_builder {
    cols_args args;
    cols_builder() { reset(); }
    void reset() { args.count = 0; args.units = 0; args.kmax = 0; args.pad = 0; }
    bool full() const { return args.count >= COLS_MAX_NODES; }
    bool empty() const { return args.count == 0; }
    bool add(int type, const uint8_t * w, int64_t row_stride, int64_t m, int64_t k, const float * x, int64_t x_stride_bytes,
             int64_t n, float * dst, ws_cursor & cur) {
        cols_node & nd = args.nodes[args.count];
        const cols_node * prev = args.count > 0 ? &args.nodes[args.count - 1] : nullptr;
        if (prev && prev->x == x && prev->k == k && prev->n == n && prev->x_stride == (int) (x_stride_bytes / 4)) {
            nd.yq = prev->yq; nd.yds = prev->yds; nd.quantize = 0;     // same activations as the previous node (gate/up, q/k/v)
        } else {
            const size_t need = cols_ws_bytes(k, n);
            if (need > cur.left) return false;
            nd.yq = reinterpret_cast<int8_t *>(cur.p);
            nd.yds = reinterpret_cast<float2 *>(cur.p + (((size_t) COLS_N * k + 127) & ~(size_t) 127));
            nd.quantize = 1; cur.p += need; cur.left -= need;
        }
        nd.w = w; nd.x = x; nd.dst = dst; nd.row_stride = row_stride; nd.m = (int) m; nd.k = (int) k; nd.n = (int) n;
        nd.x_stride = (int) (x_stride_bytes / 4); nd.unit0 = args.units; nd.type = (unsigned char) type; nd.pad[0] = nd.pad[1] = 0;
        args.units += (int) ((m + 15) / 16);
        if (nd.k > args.kmax) args.kmax = nd.k;
        args.count++;
        return true;
    }
};

static int cols_launch(cols_builder & cb, cudaStream_t st) {
    if (cb.empty()) return B200_OK;
    const dev_caps & caps = current_caps();
    cols_args & a = cb.args;
    static const int pdl = env_int("GGML_B200_FUSED_PDL", 1);
    {
        const dim3 grid((unsigned) ceil_div(a.kmax / 32, 8), (unsigned) a.count, COLS_N);
        k_quantize_cols<<<grid, 256, 0, st>>>(a);
        b200_count_launch(1);
    }
    int ctas = caps.sm_count * COLS_CTAS;
    if (ctas > a.units) ctas = a.units;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) ctas); cfg.blockDim = dim3(COLS_NW * 32);
    cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    // (whole tiles per warp, k_mmvq_cols<1>, measured 7 % slower than the K split even on long lists,
    // profiles/r01o_bs_probe.txt; that instantiation is no longer built)
    const cudaError_t e = cudaLaunchKernelEx(&cfg, k_mmvq_cols<8>, a);
    if (e != cudaSuccess) return -(int) e;
    b200_count_launch(1);
    return B200_OK;
}

} // namespace b200
