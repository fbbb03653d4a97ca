// mmq_i8.cuh — batched quantized MUL_MAT on the int8 tensor cores: the q8_1 route the north star names.
//
//   dst[m, n] = W[m, k] (Q4_0 / Q8_0 / Q4_K / Q5_K blocks)  x  X[n, k]^T (f32)
//
// Same arithmetic as the reference's mul_mat_q (src/ggml-cuda/mmq.cuh:796-931, 966-1092; quantize_mmq_q8_1,
// src/ggml-cuda/quantize.cu:50-143): activations are quantized to q8_1 (int8 + (d, d * sum) per 32), weights stay
// the integers they are, one int8 MMA per 32-wide block, and the block's two scales are applied to the int32
// result: dst += dW * dY * (qW . qY) - mW * sY.  What differs is where each part runs on B200:
//   * the MMA is tcgen05.mma.kind::i8 (SASS UTCIMMA), M = 128, N = BN, K = 32, operands in shared memory
//     (K-major, SWIZZLE_128B), every 32-block into its OWN TMEM column set (512 columns = 512 / BN sets in flight);
//   * producer warps turn raw quant slabs (2-D TMA tiles, as in mmq_tc.cuh) into int8 rows: nibbles are only
//     masked / shifted (no dequantisation, half the shared-memory bytes of the bf16 route), the per-block
//     (d * sc, dmin * m) pairs go to a small table beside the stage;
//   * activations: k_quantize_mmq_q8_1 writes an int8 plane [n][k] (fetched by 2-D TMA with the same swizzle) and
//     (d, s) pairs [k / 128][n][4] (one bulk copy per stage);
//   * eight epilogue warps drain the column sets as they complete (tcgen05.ld), convert, scale and accumulate in
//     f32 registers while the next blocks' MMAs run.
// The drain is the bound: 128 x BN int32 per 32-deep block have to cross the CUDA cores (I2F + 2..3 FMA each), about
// 4x the time the tensor pipe needs for the block's MACs (DESIGN.md §4.4 has the measured comparison with the bf16
// route of mmq_tc.cuh, which folds the scales into the operand and needs no drain).  Q6_K (scales per 16 weights:
// two scale groups per K = 32 MMA) has no int8 route and takes the bf16 kernel.
#pragma once
#include "mmq_tc.cuh"

namespace b200 {

__device__ __forceinline__ void umma_i8_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, 0, 0;\n"                                      // always a fresh accumulator
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
        "}\n" :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
}

// ================================================================= activations: f32 -> q8_1 for the tile kernel
// (the block_q8_1_mmq of this backend, a6/a8)   yq: int8 [n_pad][k]      ds: float2 (d, d * sum) [k / 128][n_pad][4]
// grid (ceil(k / 32 / 8), n_pad), block 256: one warp per 32 values, q8_1 arithmetic of quantize_row_q8_1_ref
// (src/ggml-quants.c:225-258); rows >= n are written as zeros.
static __global__ void __launch_bounds__(256) k_quantize_mmq_q8_1(const uint8_t * __restrict__ x, int64_t x_row_stride, int64_t k, int64_t n, int64_t n_pad,
                                                                   int8_t * __restrict__ yq, float2 * __restrict__ ds) {
    const int64_t row = blockIdx.y;
    const int kb = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if ((int64_t) kb * 32 >= k) {                           // the tail of the last group of four: zero scales (the tile kernel walks whole groups)
        if ((int64_t) kb < (k + 127) / 128 * 4 && lane == 0) ds[((int64_t) (kb >> 2) * n_pad + row) * 4 + (kb & 3)] = make_float2(0.0f, 0.0f);
        return;
    }
    float v = 0.0f;
    if (row < n) v = __ldg(reinterpret_cast<const float *>(x + row * x_row_stride) + kb * 32 + lane);
    const float amax = warp_max(fabsf(v));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const float r  = roundf(__fmul_rn(v, id));
    const int   qi = fabsf(r) < 2147483648.0f ? (int) r : (int) 0x80000000;
    const int   q8 = (int) (int8_t) (qi & 0xFF);
    yq[row * k + kb * 32 + lane] = (int8_t) q8;
    int sum = q8;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, off);
    if (lane == 0) {
        const float dh = __half2float(__float2half_rn(d));
        ds[((int64_t) (kb >> 2) * n_pad + row) * 4 + (kb & 3)] = make_float2(dh, dh * (float) sum);
    }
}

// ================================================================= weights: quant slab -> int8 rows + scale pairs
// A producer thread owns one weight row.  expand128<HALF> writes elements [128 HALF, 128 HALF + 128) of the 256-element
// slab as eight 16-byte chunks of int8 into the row of a SWIZZLE_128B stage and the four (dW, mW) pairs of its 32-blocks.
__device__ __forceinline__ void store_chunk_i8(uint8_t * a_row, int row7, int c, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3) {
    *reinterpret_cast<uint4 *>(a_row + ((c ^ row7) << 4)) = make_uint4(w0, w1, w2, w3);
}

template <int T> struct expander_i8;

template <bool HAS_QH> struct expander_i8_k45 {
    static constexpr bool has_min = true;
    uint32_t hdr[4];
    __device__ __forceinline__ void begin_slab(const uint8_t * s, int) { slab_words<0, 4>(s, hdr); }
    template <int HALF> __device__ __forceinline__ void expand128(const uint8_t * s, uint8_t * a_row, int row7, float2 * sc_out) {
        const float d = f16_bits_to_f32(hdr[0]), dmin = f16_bits_to_f32(hdr[0] >> 16);
        constexpr int Q0 = HAS_QH ? 12 : 4;
        uint32_t qq[16], h[8];
        slab_words<Q0 + 16 * HALF, 16>(s, qq);                 // the 64 quant bytes of groups 2 HALF and 2 HALF + 1
        if (HAS_QH) slab_words<4, 8>(s, h);
#pragma unroll
        for (int gg = 0; gg < 2; gg++) {                       // 64-element group g = 2 HALF + gg: 32 quant bytes, low / high nibbles
            const uint32_t * q = qq + 8 * gg;
            const int g = 2 * HALF + gg;
#pragma unroll
            for (int hi = 0; hi < 2; hi++) {
                int sc, mn;
                k4_scale_min(hdr[1], hdr[2], hdr[3], 2 * g + hi, sc, mn);
                sc_out[2 * gg + hi] = make_float2(d * (float) sc, dmin * (float) mn);
#pragma unroll
                for (int c = 0; c < 2; c++) {
                    uint32_t w[4];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        w[j] = (q[4 * c + j] >> (4 * hi)) & 0x0F0F0F0Fu;
                        if (HAS_QH) w[j] |= ((h[4 * c + j] >> (2 * g + hi)) & 0x01010101u) << 4;
                    }
                    store_chunk_i8(a_row, row7, 4 * gg + 2 * hi + c, w[0], w[1], w[2], w[3]);
                }
            }
        }
    }
};
template <> struct expander_i8<T_Q4_K> : expander_i8_k45<false> {};
template <> struct expander_i8<T_Q5_K> : expander_i8_k45<true> {};

// Q4_0: eight 18-byte blocks per slab (words 4.5 b ..); element j / j + 16 of a block = low / high nibble of byte j.
// The weights enter the MMA as q (0..15); the -8 goes through the activation sums: w = d q - 8 d.
template <> struct expander_i8<T_Q4_0> {
    static constexpr bool has_min = true;
    int nvalid;
    __device__ __forceinline__ void begin_slab(const uint8_t *, int nv) { nvalid = nv; }
    template <int HALF> __device__ __forceinline__ void expand128(const uint8_t * s, uint8_t * a_row, int row7, float2 * sc_out) {
        uint32_t w[18];
        slab_words<18 * HALF, 18>(s, w);
#pragma unroll
        for (int b = 0; b < 4; b++) {
            constexpr int B0[4] = { 0, 4, 9, 13 };
            uint32_t qs[4];
#pragma unroll
            for (int i = 0; i < 4; i++) qs[i] = (b & 1) == 0 ? __funnelshift_r(w[B0[b] + i], w[B0[b] + i + 1], 16) : w[B0[b] + 1 + i];
            const bool ok = 4 * HALF + b < nvalid;
            const float d = ok ? f16_bits_to_f32((b & 1) == 0 ? w[B0[b]] : (w[B0[b]] >> 16)) : 0.0f;
            sc_out[b] = make_float2(d, 8.0f * d);
            const uint32_t keep = ok ? 0x0F0F0F0Fu : 0u;
            store_chunk_i8(a_row, row7, 2 * b,     qs[0] & keep, qs[1] & keep, qs[2] & keep, qs[3] & keep);
            store_chunk_i8(a_row, row7, 2 * b + 1, (qs[0] >> 4) & keep, (qs[1] >> 4) & keep, (qs[2] >> 4) & keep, (qs[3] >> 4) & keep);
        }
    }
};

// Q8_0: eight 34-byte blocks per slab (words 8.5 b ..)
template <> struct expander_i8<T_Q8_0> {
    static constexpr bool has_min = false;
    int nvalid;
    __device__ __forceinline__ void begin_slab(const uint8_t *, int nv) { nvalid = nv; }
    template <int HALF> __device__ __forceinline__ void expand128(const uint8_t * s, uint8_t * a_row, int row7, float2 * sc_out) {
        uint32_t x[34];
        slab_words<34 * HALF, 34>(s, x);
#pragma unroll
        for (int b = 0; b < 4; b++) {
            constexpr int B0[4] = { 0, 8, 17, 25 };
            const uint32_t * w = x + B0[b];
            uint32_t q[8];
#pragma unroll
            for (int i = 0; i < 8; i++) q[i] = (b & 1) == 0 ? __funnelshift_r(w[i], w[i + 1], 16) : w[1 + i];
            const bool ok = 4 * HALF + b < nvalid;
            const float d = ok ? f16_bits_to_f32((b & 1) == 0 ? w[0] : (w[0] >> 16)) : 0.0f;
            sc_out[b] = make_float2(d, 0.0f);
            const uint32_t keep = ok ? 0xFFFFFFFFu : 0u;
            store_chunk_i8(a_row, row7, 2 * b,     q[0] & keep, q[1] & keep, q[2] & keep, q[3] & keep);
            store_chunk_i8(a_row, row7, 2 * b + 1, q[4] & keep, q[5] & keep, q[6] & keep, q[7] & keep);
        }
    }
};

// ================================================================= the kernel
struct mmq_i8_args {
    int64_t m, k, n, n_pad;
    float * dst;               // dst[j * dst_col_stride + i]
    int64_t dst_col_stride;    // floats
    const float2 * ds;         // activation (d, s) pairs [k / 128][n_pad][4]
    int     nblk;              // weight blocks per row (k / qk)
    int     raw_pitch;
};

constexpr int MMQI_BM = 128;
constexpr int MMQI_ST = 3;            // stages of 128 K elements (128 bytes per row)
constexpr int MMQI_RS = 3;            // raw slab ring
constexpr int MMQI_PROD = 4, MMQI_EPI = 8;
constexpr int MMQI_THREADS = (MMQI_PROD + MMQI_EPI + 2) * 32;

template <int BN> struct mmq_i8_smem {
    static constexpr int a_bytes = MMQI_BM * 128, b_bytes = BN * 128;
    static constexpr int sa_bytes = MMQI_BM * 4 * 8, sy_bytes = BN * 4 * 8;       // (dW, mW) and (dY, sY) per 32-block
    static constexpr int a_off = 0;
    static constexpr int b_off = a_off + MMQI_ST * a_bytes;
    static constexpr int sa_off = b_off + MMQI_ST * b_bytes;
    static constexpr int sy_off = sa_off + MMQI_ST * sa_bytes;
    static constexpr int raw_off = sy_off + MMQI_ST * sy_bytes;
    static size_t total(int raw_pitch, int rs) { return (size_t) raw_off + (size_t) rs * MMQI_BM * raw_pitch + 512 + 1024; }
};

template <int T, int BN>
__global__ void __launch_bounds__(MMQI_THREADS, 1)
k_mmq_i8(const mmq_i8_args a, const __grid_constant__ CUtensorMap tmap_b, const __grid_constant__ CUtensorMap tmap_w) {
    using SM = mmq_i8_smem<BN>;
    constexpr int ST = MMQI_ST, RS = MMQI_RS;
    constexpr int BLK_PER_SLAB = 256 / fmt<T>::qk;
    constexpr int NBUF = 512 / BN > 8 ? 8 : 512 / BN;          // TMEM column sets (one 32-block each) in flight
    constexpr int CPT = BN / 2;                                // columns an epilogue thread owns

    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    uint8_t * a_tiles = smem + SM::a_off;
    uint8_t * b_tiles = smem + SM::b_off;
    float2  * sA = reinterpret_cast<float2 *>(smem + SM::sa_off);       // [ST][128][4]
    float2  * sY = reinterpret_cast<float2 *>(smem + SM::sy_off);       // [ST][BN][4]
    uint8_t * raw = smem + SM::raw_off;
    uint8_t * ctl = raw + (size_t) RS * MMQI_BM * a.raw_pitch;
    uint64_t * full       = reinterpret_cast<uint64_t *>(ctl);           // [ST]   A + scales written, B + activation scales landed
    uint64_t * empty      = full + ST;                                   // [ST]   MMAs retired and the epilogue done with the scales
    uint64_t * raw_full   = empty + ST;                                  // [RS]
    uint64_t * raw_empty  = raw_full + RS;                               // [RS]
    uint64_t * tmem_full  = raw_empty + RS;                              // [8]    the block's MMA has completed
    uint64_t * tmem_empty = tmem_full + 8;                               // [8]    the epilogue has drained the column set
    uint32_t * tmem_slot  = reinterpret_cast<uint32_t *>(tmem_empty + 8);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t m0 = (int64_t) blockIdx.x * MMQI_BM;
    const int64_t n0 = (int64_t) blockIdx.y * BN;
    const int nkt   = (int) ((a.k + 127) / 128);                         // stages
    const int nslab = (int) ((a.k + 255) / 256);
    const int nb    = nkt * 4;                                           // 32-blocks (a k tail is zero-filled by the producers)

    if (tid == 0) {
        for (int s = 0; s < ST; s++) { mbar_init(&full[s], MMQI_PROD * 32 + 1); mbar_init(&empty[s], 1 + MMQI_EPI); }
        for (int s = 0; s < RS; s++) { mbar_init(&raw_full[s], 1); mbar_init(&raw_empty[s], MMQI_PROD * 32); }
        for (int s = 0; s < 8; s++) { mbar_init(&tmem_full[s], 1); mbar_init(&tmem_empty[s], MMQI_EPI); }
        mbar_fence_init();
    }
    if (warp == MMQI_PROD + MMQI_EPI + 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < MMQI_PROD) {
        // ===================================================== weight producers: one thread per row
        const int prow = tid;
        const int row7 = prow & 7;
        auto slab_valid_blocks = [&](int s) {
            const int left = a.nblk - s * BLK_PER_SLAB;
            return left < BLK_PER_SLAB ? left : BLK_PER_SLAB;
        };
        expander_i8<T> ex;
        for (int s = 0; s < nslab; s++) {
            const int r = s % RS;
            mbar_wait(&raw_full[r], (uint32_t) ((s / RS) & 1));
            const uint8_t * slab_ptr = raw + ((size_t) r * MMQI_BM + prow) * a.raw_pitch;
            ex.begin_slab(slab_ptr, slab_valid_blocks(s));
#pragma unroll
            for (int half = 0; half < 2; half++) {
                const int kt = s * 2 + half;
                if (kt < nkt) {
                    const int st = kt % ST;
                    if (kt >= ST) mbar_wait(&empty[st], (uint32_t) (((kt / ST) - 1) & 1));
                    uint8_t * a_row = a_tiles + (size_t) st * SM::a_bytes + prow * 128;
                    float2 * sc = sA + ((size_t) st * MMQI_BM + prow) * 4;
                    if (half == 0) ex.template expand128<0>(slab_ptr, a_row, row7, sc);
                    else           ex.template expand128<1>(slab_ptr, a_row, row7, sc);
                    fence_proxy_async_smem();
                    mbar_arrive(&full[st]);
                }
            }
            mbar_arrive(&raw_empty[r]);
        }
    } else if (warp < MMQI_PROD + MMQI_EPI) {
        // ===================================================== epilogue: drain each block's column set, scale, accumulate
        const int ew = warp - MMQI_PROD;
        const int quarter = warp & 3;                    // a warp reaches TMEM lanes 32 (warp % 4) .. + 31
        const int half = ew >> 2;                        // which half of the tile's columns (warps 4..7 / 8..11: warp % 4 covers 0..3 in both)
        const int row = quarter * 32 + lane;
        float acc[CPT];
#pragma unroll
        for (int j = 0; j < CPT; j++) acc[j] = 0.0f;
        for (int b = 0; b < nb; b++) {
            const int buf = b % NBUF, st = (b >> 2) % ST, k4 = b & 3;
            mbar_wait(&tmem_full[buf], (uint32_t) ((b / NBUF) & 1));
            tc_fence_after();
            const float2 wA = sA[((size_t) st * MMQI_BM + row) * 4 + k4];
            const float2 * yp = sY + ((size_t) st * BN + half * CPT) * 4 + k4;
            const uint32_t taddr = tmem_base + ((uint32_t) (quarter * 32) << 16) + (uint32_t) (buf * BN + half * CPT);
#pragma unroll
            for (int c0 = 0; c0 < CPT; c0 += 16) {
                uint32_t v[16];
                tmem_ld16(taddr + (uint32_t) c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    const float2 y = yp[(size_t) (c0 + j) * 4];                 // (dY, sY) of column c0 + j: the same address for the whole warp
                    float t = acc[c0 + j];
                    t = fmaf((float) (int) v[j], wA.x * y.x, t);
                    if (expander_i8<T>::has_min) t = fmaf(-wA.y, y.y, t);
                    acc[c0 + j] = t;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&tmem_empty[buf]);
                if (k4 == 3) mbar_arrive(&empty[st]);            // done with the stage's scale tables
            }
        }
        const int64_t out_row = m0 + row;
        if (out_row < a.m) {
#pragma unroll
            for (int j = 0; j < CPT; j++) {
                const int64_t col = n0 + half * CPT + j;
                if (col < a.n) a.dst[col * a.dst_col_stride + out_row] = acc[j];
            }
        }
    } else if (warp == MMQI_PROD + MMQI_EPI) {
        // ===================================================== loader: raw weight slabs, activation tiles and scales
        if (lane == 0) {
            int next_slab = 0;
            for (int kt = 0; kt < nkt; kt++) {
                while (next_slab < nslab && next_slab <= kt / 2 + RS - 1) {
                    const int r = next_slab % RS;
                    if (next_slab >= RS) mbar_wait(&raw_empty[r], (uint32_t) (((next_slab / RS) - 1) & 1));
                    mbar_arrive_expect_tx(&raw_full[r], (uint32_t) (MMQI_BM * slab<T>::bytes));
                    tma_load_2d(raw + (size_t) r * MMQI_BM * a.raw_pitch, &tmap_w, next_slab * (slab<T>::bytes / 4), (int) m0, &raw_full[r]);
                    next_slab++;
                }
                const int st = kt % ST;
                if (kt >= ST) mbar_wait(&empty[st], (uint32_t) (((kt / ST) - 1) & 1));
                mbar_arrive_expect_tx(&full[st], (uint32_t) (SM::b_bytes + SM::sy_bytes));
                tma_load_2d(b_tiles + (size_t) st * SM::b_bytes, &tmap_b, kt * 128, (int) n0, &full[st]);
                bulk_g2s(sY + (size_t) st * BN * 4, a.ds + ((int64_t) kt * a.n_pad + n0) * 4, (uint32_t) SM::sy_bytes, &full[st]);
            }
        }
    } else {
        // ===================================================== MMA issuer: one K = 32 MMA per block, each into its own column set
        if (lane == 0) {
            // D = s32 (2), A = s8 (1), B = s8 (1), both K-major, N = BN, M = 128  (weights 0..31 and -128..127 are both valid s8)
            constexpr uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t) (BN >> 3) << 17) | ((uint32_t) (MMQI_BM >> 4) << 24);
            for (int kt = 0; kt < nkt; kt++) {
                const int st = kt % ST;
                mbar_wait(&full[st], (uint32_t) ((kt / ST) & 1));
                tc_fence_after();
                const uint64_t da = umma_desc_sw128(a_tiles + (size_t) st * SM::a_bytes);
                const uint64_t db = umma_desc_sw128(b_tiles + (size_t) st * SM::b_bytes);
#pragma unroll
                for (int k4 = 0; k4 < 4; k4++) {
                    const int b = kt * 4 + k4, buf = b % NBUF;
                    if (b >= NBUF) { mbar_wait(&tmem_empty[buf], (uint32_t) (((b / NBUF) - 1) & 1)); tc_fence_after(); }
                    umma_i8_ss(tmem_base + (uint32_t) (buf * BN), da + (uint64_t) (2 * k4), db + (uint64_t) (2 * k4), idesc);
                    umma_commit(&tmem_full[buf]);
                }
                umma_commit(&empty[st]);                 // the stage's operands are free when these MMAs retire
            }
        }
    }

    __syncthreads();
    if (warp == MMQI_PROD + MMQI_EPI + 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

} // namespace b200
