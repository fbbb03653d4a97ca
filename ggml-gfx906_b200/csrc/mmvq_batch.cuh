// mmvq_batch.cuh — the persistent decode mat-vec: a whole LIST of independent n = 1 MUL_MATs in ONE launch.
//
//   for every node i:   dst_i[m_i] = W_i[m_i, k_i] (ggml quant blocks) . x_i[k_i] (f32)
//
// Why: a decode mat-vec is tiny next to a B200 (4096 x 4096 Q4_K = 9.4 MB = 1.3 us of HBM time, one
// DRAM round trip of bytes per SM), so one launch per MUL_MAT (what the reference's graph loop does,
// src/ggml-cuda/ggml-cuda.cu:2870-2960 -> mmvq.cu:500-570, plus a quantize_q8_1 launch each) spends most
// of its time filling and draining the machine: measured 8 us per 4096^2 mat-vec = 1.2 TB/s
// (profiles/r01b_mmvq_probe_per_launch.txt) against a 7.2 TB/s read ceiling for a TMA ring
// (profiles/r01b_stream_probe.txt).  Here the machine is filled once per list:
//
//   * k_quantize_batch: ONE small launch turns every node's x into q8_1 values (bit-exact
//     quantize_row_q8_1_ref, src/ggml-quants.c:225-258) in the caller's workspace: int8 plane [k] followed
//     by the fp16-rounded scales as f32 [k/32] — 1.125 bytes per activation, nodes sharing x share it
//     (the int8 plane is bank-swizzled per 128-value item for the consumers' register loads);
//   * k_mmvq_batch: one CTA per SM, resident for the whole list; every node's rows are dealt to the CTAs as
//     contiguous slabs (rotated per node so that short last slabs do not always hit the same SM);
//       warp 0 (weight producer) streams the CTA's slabs of ALL nodes, back to back, through one ring of
//         shared-memory stages with bulk async copies (cp.async.bulk, SASS UBLKCP) on full/empty
//         mbarriers: the weights of node i+1.. are in flight while node i is reduced, HBM never idles;
//       warp 1 (activation producer) bulk-copies the quantized activations of the next nodes into a small
//         ring of activation buffers (ready/free mbarriers);
//       14 consumer warps: a lane owns 128 consecutive weights of a row (hitem<T>, mmvq_fused.cuh) and keeps
//         the matching q8 activations in registers for the node; W consecutive warps cover a row, the
//         G = 14 / W groups (12 when W = 1) take the row pairs of the stages round-robin; partials of a row meet in shared
//         memory in fixed order, so a row's value never depends on the grid, the list or the GPU that
//         computes it (row-split results stay bit-identical).  Every consumer warp observes every stage
//         (mbarrier phases are parity bits: a waiter that skipped stages could mistake phase n-2 for n).
//
// Dependent nodes (a decode pass is a chain: wo reads what q/k/v wrote, gate/up read wo ...) ride in the SAME launch:
// a node that reads (or overwrites) what an earlier node of the launch writes carries that node's index as `dep`.
//   * every CTA bumps a per-node completion counter in global memory (release, gpu scope) when its consumer warps
//     have stored their last row of a node that something depends on;
//   * the activation-producer warp of a dependent node acquire-spins on done[dep] == gridDim.x before it publishes
//     the node to the consumers, which then quantize x = the producer node's dst (L2 -> registers -> the
//     activation buffer, the same q8_1 values and layout as k_quantize_batch) themselves;
//   * weights keep streaming through the ring meanwhile, so the cost of a chain link is one L2 round trip for the
//     flag, one for x and the quantization (~1 us), not a kernel boundary (7-17 us measured, profiles/r01m).
// The grid is one CTA per SM with 226 KB of shared memory each: all CTAs are co-resident, so the spins terminate.
// The role CUDA-graph capture plays in the reference (src/ggml-cuda/ggml-cuda.cu:2876-3041) is played by this.
//
// Row split over the GPUs of a box (reference: peer-copy gather after every MUL_MAT, ggml-cuda.cu:1670-1739): the
// gather is FUSED into this kernel.  Every rank runs the same list on its row blocks; a node whose dst lies in the
// symmetric region (one copy per rank, peer-mapped: NVLink / NVSwitch) stores each finished row into EVERY rank's copy
// (plain st.global on the peer addresses, as the rows are produced); the CTAs of a rank meet on a local arrival counter
// (system-scope fence, then a gpu-scope add) and the last of them bumps the node's completion counter on every rank
// (one arrival per rank).  The counters are double-buffered by the parity of the launch count ("epoch", bumped by the
// launch's quantize pre-pass, which also zeroes the OTHER set: no rank can touch that set before this rank has finished
// the current launch, because every rank's closing barrier needs it).  A dependent node waits for the gathered vector exactly as it waits for
// a local one, and CTA 0 leaves the kernel only when the last scattered node is complete on all ranks: when the
// launch is done, every rank holds every row.
// b200_mul_mat_batch derives `dep` from the pointer ranges; hazards between different launches still cut the list.
#pragma once
#include <vector>
#include "mmvq_fused.cuh"

namespace b200 {

constexpr int BATCH_MAX_NODES  = 448;
constexpr int BATCH_NWC        = 14;     // consumer warps at most; a launch runs 12 or 14 (batch_args::nwc)
constexpr int BATCH_AHDR       = 128;    // bytes in front of each activation buffer: the node as the consumers need it
constexpr int BATCH_THREADS    = (2 + BATCH_NWC) * 32;
constexpr int BATCH_MAX_STAGES = 16;
constexpr int BATCH_MAX_ABUF   = 4;      // activation buffers
constexpr int BATCH_RMAX       = 16;     // rows per stage, at most, when a row needs several warps (partials buffer)
constexpr int BATCH_RMAX_W1    = 96;     // rows per stage, at most, otherwise
constexpr int BATCH_ROT        = 61;     // slab rotation per node (coprime with the SM count)

struct batch_node {                      // 64 bytes
    const uint8_t * w;                   // first weight row
    const float   * x;                   // k floats, 16-byte aligned
    float         * dst;                 // m floats
    const uint8_t * acts;                // quantized activations in the workspace: int8[k] | f32 d[k/32]
    int64_t         row_stride;          // bytes between rows
    int   m, k;
    int   row_bytes;
    int   rows_per_cta;                  // slab height (even)
    unsigned char type, R, W, G, contiguous, qflags;             // qflags: BATCH_Q_*
    short dep;                                                   // index of the node of this launch to wait for, -1: none
};
constexpr int BATCH_Q_PRE    = 1;    // k_quantize_batch writes this node's q8 record (nd.acts) before the launch
constexpr int BATCH_Q_LOAD   = 2;    // the activation producer bulk-copies nd.acts into the activation buffer
constexpr int BATCH_Q_INK    = 4;    // the consumers quantize nd.x (written by node `dep` of this launch) into the buffer
constexpr int BATCH_Q_SAME   = 8;    // same x, k and format as the previous node: the consumers keep their registers
constexpr int BATCH_Q_SIGNAL = 16;   // a later node waits for this one: CTAs bump done[i] after their last row
constexpr int BATCH_Q_SCATTER = 32;  // dst lies in the symmetric region: rows and the completion counter go to every rank
constexpr int BATCH_MAX_PEERS = 7;   // other ranks of the box
static_assert(sizeof(batch_node) == 64, "batch_node is part of the kernel parameter block");

struct batch_args {
    int count, S, slot_stride, acts_stride;
    int off_part, off_acts, off_slots, const_weights;
    int NB, kmax, nwc, pad;
    unsigned * done;                     // per-node completion counters (workspace, zeroed by k_quantize_batch; or the symmetric
                                         // region's control block, never zeroed: see epoch)
    unsigned * epoch;                    // row split: launches of this family so far (bumped by k_quantize_batch); else NULL
    long long  peer_delta[BATCH_MAX_PEERS];   // row split: byte distance from the local copy of the symmetric region to rank p's
    int        n_peers, last_scatter;    // other ranks; index of the last BATCH_Q_SCATTER node (the launch's closing barrier), -1
#ifdef B200_TUNING
    unsigned long long * dbg;            // [node < 96][CTA][4] globaltimer stamps: consumers released, CTA done, flag seen, first stage
#endif
    batch_node nodes[BATCH_MAX_NODES];
};
#ifdef B200_TUNING
#define BATCH_STAMP(node, slot) do { if (a.dbg && (node) < 96) a.dbg[((size_t) (node) * gridDim.x + blockIdx.x) * 8 + (slot)] = gtime(); } while (0)
#else
#define BATCH_STAMP(node, slot) do { } while (0)
#endif
#ifdef B200_TUNING
// cycle accumulators of one thread (row 95 - which of the stamp table, 8 per CTA): T0 = clock64() .. ACC(var) adds the cycles since
#define BATCH_ACC_DECL(var) long long var = 0
#define BATCH_ACC_T0(t) const long long t = clock64()
#define BATCH_ACC_ADD(var, t) var += clock64() - t
#define BATCH_ACC_OUT(row, slot, var) do { if (a.dbg) a.dbg[((size_t) (row) * gridDim.x + blockIdx.x) * 8 + (slot)] = (unsigned long long) (var); } while (0)
#define BATCH_TUN_PARAMS , long long & acc_full, long long & acc_bar
#define BATCH_TUN_ARGS , acc_full, acc_bar
#else
#define BATCH_TUN_PARAMS
#define BATCH_TUN_ARGS
#define BATCH_ACC_DECL(var) do { } while (0)
#define BATCH_ACC_T0(t) do { } while (0)
#define BATCH_ACC_ADD(var, t) do { } while (0)
#define BATCH_ACC_OUT(row, slot, var) do { } while (0)
#endif

// gpu-scope release / acquire on the completion counters
__device__ __forceinline__ void red_release_gpu_add(unsigned * p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_relaxed_gpu(const unsigned * p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_sys_add(unsigned * p, unsigned v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_relaxed_sys_add(unsigned * p, unsigned v) {
    asm volatile("red.relaxed.sys.global.add.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_relaxed_sys(const unsigned * p) {
    unsigned v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// wait until counter *ctr has reached `target` (counters only grow; wrap-safe).  One lane polls; a bound turns a logic
// error (or a rank that never launched) into a trap instead of a hang.
__device__ __forceinline__ void batch_wait_counter(const unsigned * ctr, unsigned target, bool sys) {
    unsigned spins = 0;
    if (sys) {
        while ((int) (ld_relaxed_sys(ctr) - target) < 0) { if (++spins > (1u << 26)) __trap(); }
        asm volatile("fence.acq_rel.sys;" ::: "memory");
    } else {
        while ((int) (ld_relaxed_gpu(ctr) - target) < 0) { if (++spins > (1u << 25)) __trap(); }
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
}

// mbarrier ops on precomputed shared-window addresses (no generic -> shared conversion per call)
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "B200A_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra B200A_DONE;\n"
        "bra B200A_WAIT;\n"
        "B200A_DONE:\n"
        "}\n" :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}

__device__ __forceinline__ int batch_row_pitch(int row_bytes) { return (row_bytes + 15 + 15 + 16) & ~15; }
__device__ __forceinline__ uint32_t batch_acts_bytes(int k) { return (uint32_t) (k + (k >> 5) * 4); }

// A node descriptor in registers.  The parameter block (up to 28 KB) is far larger than the constant
// cache: reading a descriptor where it is needed costs a chain of cache misses per node and per warp
// (measured: 23 % of the kernel).  Instead every lane of a warp keeps ONE descriptor (node base + lane),
// fetched 32 nodes at a time, and the warp broadcasts the current node's fields by shuffle.
struct node_r {
    const uint8_t * w; const float * x; float * dst; const uint8_t * acts; int64_t row_stride;
    int m, k, row_bytes, rows_per_cta, type, R, W, G, contiguous, qflags, dep;
};
__device__ __forceinline__ uint64_t shfl_u64(uint64_t v, int src) {
    const uint32_t lo = __shfl_sync(0xFFFFFFFFu, (uint32_t) v, src), hi = __shfl_sync(0xFFFFFFFFu, (uint32_t) (v >> 32), src);
    return ((uint64_t) hi << 32) | lo;
}
struct node_cache {
    node_r c;
    int base;
    __device__ __forceinline__ node_cache() : base(-64) {}
    __device__ __forceinline__ node_r get(const batch_args & a, int i, int lane) {     // convergent: all 32 lanes call
        if (i >= base + 32) {
            base = i & ~31;
            const int j = base + lane < a.count ? base + lane : a.count - 1;
            const batch_node & nd = a.nodes[j];
            c.w = nd.w; c.x = nd.x; c.dst = nd.dst; c.acts = nd.acts; c.row_stride = nd.row_stride;
            c.m = nd.m; c.k = nd.k; c.row_bytes = nd.row_bytes; c.rows_per_cta = nd.rows_per_cta;
            c.type = nd.type; c.R = nd.R; c.W = nd.W; c.G = nd.G; c.contiguous = nd.contiguous;
            c.qflags = nd.qflags; c.dep = nd.dep;
        }
        const int src = i & 31;
        node_r r;
        r.w = reinterpret_cast<const uint8_t *>(shfl_u64(reinterpret_cast<uint64_t>(c.w), src));
        r.x = reinterpret_cast<const float *>(shfl_u64(reinterpret_cast<uint64_t>(c.x), src));
        r.dst = reinterpret_cast<float *>(shfl_u64(reinterpret_cast<uint64_t>(c.dst), src));
        r.acts = reinterpret_cast<const uint8_t *>(shfl_u64(reinterpret_cast<uint64_t>(c.acts), src));
        r.row_stride = (int64_t) shfl_u64((uint64_t) c.row_stride, src);
        r.m = __shfl_sync(0xFFFFFFFFu, c.m, src); r.k = __shfl_sync(0xFFFFFFFFu, c.k, src);
        r.row_bytes = __shfl_sync(0xFFFFFFFFu, c.row_bytes, src); r.rows_per_cta = __shfl_sync(0xFFFFFFFFu, c.rows_per_cta, src);
        const int packed = __shfl_sync(0xFFFFFFFFu, c.type | (c.R << 8) | (c.W << 16) | (c.G << 20) | (c.contiguous << 24), src);
        r.type = packed & 0xFF; r.R = (packed >> 8) & 0xFF; r.W = (packed >> 16) & 0xF; r.G = (packed >> 20) & 0xF; r.contiguous = (packed >> 24) & 1;
        const int packed2 = __shfl_sync(0xFFFFFFFFu, c.qflags | (c.dep << 8), src);
        r.qflags = packed2 & 0xFF; r.dep = packed2 >> 8;              // arithmetic shift: dep = -1 survives
        return r;
    }
};

// v: this CTA's slab index for the node = (blockIdx.x + i * BATCH_ROT) % gridDim.x, kept incrementally
__device__ __forceinline__ void batch_slab(const node_r & nd, int & v, int & r_begin, int & nrows) {
    const int64_t rb = (int64_t) v * nd.rows_per_cta;
    int64_t n = (int64_t) nd.m - rb;
    if (n > nd.rows_per_cta) n = nd.rows_per_cta;
    if (n < 0) n = 0;
    r_begin = (int) (rb < nd.m ? rb : 0);
    nrows = (int) n;
    v += BATCH_ROT % (int) gridDim.x;
    if (v >= (int) gridDim.x) v -= (int) gridDim.x;
}

// ---------------------------------------------------------------- all activations of a list: f32 -> q8
// grid (ceil(kmax / 128 / 8), count), block 256: one warp per 128 values -- 8 lanes per 32-value block, 4 consecutive values per
// lane (one 16-byte load), the arithmetic of k_quantize_acts.  (One warp per 32 values was 9 675 CTAs per Llama-2-7B list, half
// of them past the end of a k = 4096 node: 9 us in front of every launch, every rank of a row split pays all of it.)
__global__ void __launch_bounds__(256) k_quantize_batch(const __grid_constant__ batch_args a) {
    const batch_node & nd = a.nodes[blockIdx.y];
    const int item = blockIdx.x * 8 + (threadIdx.x >> 5);           // 128 values
    const int lane = threadIdx.x & 31;
    if (a.epoch) {
        // row split: this launch counts in set (epoch & 1); the other set is zeroed for the next launch (see the header)
        if (blockIdx.x == 0 && blockIdx.y == 0) {
            __shared__ unsigned e_new;
            if (threadIdx.x == 0) { e_new = *a.epoch + 1u; *a.epoch = e_new; }
            __syncthreads();
            unsigned * other = a.done + ((e_new + 1u) & 1u) * BATCH_MAX_NODES;
            for (int j = threadIdx.x; j < BATCH_MAX_NODES; j += 256) { other[j] = 0u; other[j + 2 * BATCH_MAX_NODES] = 0u; }   // + the local arrivals
        }
    } else if (blockIdx.x == 0 && threadIdx.x == 0 && a.done) {
        a.done[blockIdx.y] = 0u;
    }
    if (!(nd.qflags & BATCH_Q_PRE) || item * 128 >= nd.k) return;    // k is a multiple of 256
    const float4 x = __ldg(reinterpret_cast<const float4 *>(nd.x + item * 128) + lane);
    float amax = fmaxf(fmaxf(fabsf(x.x), fabsf(x.y)), fmaxf(fabsf(x.z), fabsf(x.w)));
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, o));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    // float -> int8 as quantize.cu does it (cvttss2si semantics of the reference's x86 build)
    auto q8 = [&](float xv) {
        const float v = roundf(__fmul_rn(xv, id));
        return (uint32_t) ((fabsf(v) < 2147483648.0f ? (int) v : (int) 0x80000000) & 0xFF);
    };
    const uint32_t pk = q8(x.x) | (q8(x.y) << 8) | (q8(x.z) << 16) | (q8(x.w) << 24);
    int8_t * q = reinterpret_cast<int8_t *>(const_cast<uint8_t *>(nd.acts));
    // 16-byte piece p of 128-value item idx is stored at piece (p + idx) % 8 of the item: the consumer lanes
    // (one item each, 128 B apart) then read piece p from eight different bank groups
    const int pc = lane >> 2;                                        // elements 4 lane .. 4 lane + 3 of the item
    *reinterpret_cast<uint32_t *>(q + item * 128 + (((pc + item) & 7) << 4) + (lane & 3) * 4) = pk;
    if ((lane & 7) == 0) reinterpret_cast<float *>(q + nd.k)[item * 4 + (lane >> 3)] = __half2float(__float2half_rn(d));
}

// What the consumers need of a node; written to the head of the activation buffer by the activation
// producer (one thread, ahead of time), read from shared memory by every consumer warp.
struct cnode {
    float * dst;
    int k, pitch;             // pitch: bytes between rows inside a stage
    int r_begin, nrows;       // this CTA's slab
    int type, R, W, G;
    int skew0, rs16;          // formats whose rows are not 16-byte aligned: (address of slab row i) % 16 = (skew0 + i * rs16) % 16
    int contiguous;
    int qflags;               // BATCH_Q_*
    const float * x;          // BATCH_Q_INK: the f32 activations to quantize (another node's dst)
    unsigned * signal;        // BATCH_Q_SIGNAL: this node's completion counter
};
static_assert(sizeof(cnode) <= BATCH_AHDR, "consumer header");

// ---------------------------------------------------------------- consumer: one node
// cc: running count of row pairs ("chunks") this CTA has seen; chunk j belongs to group j % G.
// The consumers of a dependent node quantize its activations themselves: x (another node's dst, complete and visible:
// the activation producer acquired its counter before publishing this node) -> q8_1 values in the activation buffer,
// bit for bit what k_quantize_batch writes (same arithmetic, same swizzle).  8 lanes per 32-value block, all consumer
// warps; x is read past L1 (it was written by other SMs during this launch).
__device__ __forceinline__ void batch_quantize_ink(const float * x, const int k, uint8_t * actbuf, const int cw, const int lane, const int nwc) {
    int8_t * q = reinterpret_cast<int8_t *>(actbuf);
    float  * yd = reinterpret_cast<float *>(actbuf + k);
    const int nblk32 = k >> 5;
    const int sub = lane >> 3, l8 = lane & 7;
    constexpr int U = 4;                                      // loads in flight per lane (8 would cover k = 11008 in one round trip but spills the consumers)
    for (int blk0 = cw * 4 + sub; blk0 < nblk32; blk0 += nwc * 4 * U) {
        float4 vv[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int blk = blk0 + u * nwc * 4;
            if (blk < nblk32) vv[u] = __ldcg(reinterpret_cast<const float4 *>(x + blk * 32 + l8 * 4));
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int blk = blk0 + u * nwc * 4;
            if (blk < nblk32) {                               // (k % 128 == 0: uniform over each group of 8 lanes)
                const float4 v = vv[u];
                float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, o));
                const float d  = __fdiv_rn(amax, 127.0f);
                const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
                const float xs[4] = { v.x, v.y, v.z, v.w };
                uint32_t packed = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const float r  = roundf(__fmul_rn(xs[j], id));
                    const int   qi = fabsf(r) < 2147483648.0f ? (int) r : (int) 0x80000000;
                    packed |= (uint32_t) (qi & 0xFF) << (8 * j);
                }
                const int e = blk * 32 + l8 * 4, idx = e >> 7, pc = (e >> 4) & 7;
                *reinterpret_cast<uint32_t *>(q + idx * 128 + (((pc + idx) & 7) << 4) + (e & 15)) = packed;
                if (l8 == 0) yd[blk] = __half2float(__float2half_rn(d));
            }
        }
    }
}

// a finished row value -> dst, and (row split) -> the same place in every other rank's copy of the symmetric region
__device__ __forceinline__ void batch_store_row(float * p, float v, const batch_args & a, bool scatter) {
    *p = v;
    if (scatter) {
#pragma unroll 1
        for (int r = 0; r < a.n_peers; r++) *reinterpret_cast<float *>(reinterpret_cast<char *>(p) + a.peer_delta[r]) = v;
    }
}

template <int T>
__device__ __forceinline__ void batch_consume(const batch_args & a, const cnode & nd, acts128 & acts, const int cw, const int lane,
                                              int & st_s, uint32_t & st_ph, int & cc, int & git,
                                              const uint8_t * actbuf, const uint32_t acts_free, const uint32_t full, const uint32_t empty,
                                              float * part, const uint8_t * slots, const int S, const int slot_stride BATCH_TUN_PARAMS) {
    using IT = hitem<T>;
    const int W = nd.W, G = nd.G, R = nd.R;
    const int r_begin = nd.r_begin, nrows = nd.nrows, pitch = nd.pitch;
    const int nbatch = (nrows + R - 1) / R;
    float * const dst = nd.dst + r_begin;
    const int g = cw / W, wg = cw - g * W;
    const int items = nd.k >> 7;
    const int idx = wg * 32 + lane;                      // the 128-weight item of a row this lane owns
    const bool active = g < G && idx < items;
    const bool scatter = (nd.qflags & BATCH_Q_SCATTER) != 0;

    // the slice of K this lane owns, in registers for the node (kept from the previous node when it has the same
    // activations, k and format: q/k/v, gate/up)
    if (active && !(nd.qflags & BATCH_Q_SAME)) {
        const int8_t * yq = reinterpret_cast<const int8_t *>(actbuf);
        const float  * yd = reinterpret_cast<const float *>(actbuf + nd.k);
        const float4 d4 = *reinterpret_cast<const float4 *>(yd + idx * 4);
        const float dd[4] = { d4.x, d4.y, d4.z, d4.w };
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int4 lo = *reinterpret_cast<const int4 *>(yq + idx * 128 + (((2 * j + idx) & 7) << 4));       // swizzled by k_quantize_batch
            const int4 hi = *reinterpret_cast<const int4 *>(yq + idx * 128 + (((2 * j + 1 + idx) & 7) << 4));
            acts.q[j][0] = lo.x; acts.q[j][1] = lo.y; acts.q[j][2] = lo.z; acts.q[j][3] = lo.w;
            acts.q[j][4] = hi.x; acts.q[j][5] = hi.y; acts.q[j][6] = hi.z; acts.q[j][7] = hi.w;
            int sum = 0, sum_hi = 0;
#pragma unroll
            for (int u = 0; u < 4; u++) { sum = __dp4a(0x01010101, acts.q[j][u], sum); sum_hi = __dp4a(0x01010101, acts.q[j][u + 4], sum_hi); }
            if (IT::need_ys16) { acts.ys16[2 * j] = sum; acts.ys16[2 * j + 1] = sum_hi; }
            sum += sum_hi;
            const float d = dd[j];
            acts.sd[j] = d * (float) sum;
            acts.d[j]  = (IT::hi_div16 && (j & 1)) ? d * 0.0625f : d;
        }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive_a(acts_free);              // this warp no longer reads the activation buffer

    // git: stages in which this group had rows (W > 1): selects the partials buffer
    int cmod = cc % G;                                    // owner group of the next chunk
    for (int lb = 0; lb < nbatch; lb++) {
        const int s = st_s;
        { BATCH_ACC_T0(t_w);
        mbar_wait_a(full + 8 * s, st_ph);                 // every lane polls (a single polling lane sleeps out its time slice)
        BATCH_ACC_ADD(acc_full, t_w); }
        if (++st_s == S) { st_s = 0; st_ph ^= 1u; }
        const int rows_here = (nrows - lb * R < R) ? (nrows - lb * R) : R;
        const int nchunks = (rows_here + FUSED_RB - 1) / FUSED_RB;
        const int c_first = g >= cmod ? g - cmod : g - cmod + G;   // first chunk of this stage that belongs to group g
        cc += nchunks;
        cmod += nchunks;
        while (cmod >= G) cmod -= G;
        float * pb = part + (size_t) ((g * 2 + (git & 1)) * W) * BATCH_RMAX;
        if (g < G && c_first < nchunks) {
            const uint8_t * slot = slots + s * slot_stride;       // 16-byte aligned formats: no skew inside a stage
            for (int c = c_first; c < nchunks; c += G) {
                const int r0 = c * FUSED_RB;
                float acc[FUSED_RB];
#pragma unroll
                for (int rr = 0; rr < FUSED_RB; rr++) acc[rr] = 0.0f;
                if (active) {
                    auto row_off = [&](int rr) -> int {
                        const int r = (r0 + rr < rows_here) ? (r0 + rr) : (rows_here - 1);   // odd tail: recomputed, never stored
                        int off = r * pitch;
                        if (!IT::aligned16) off += nd.contiguous ? ((nd.skew0 + lb * R * nd.rs16) & 15) : ((nd.skew0 + (lb * R + r) * nd.rs16) & 15);
                        return off;
                    };
                    if (IT::aligned16) {
                        // light items: both rows' words in flight before the first dot
                        typename IT::wts w[FUSED_RB];
#pragma unroll
                        for (int rr = 0; rr < FUSED_RB; rr++) IT::load(w[rr], slot + row_off(rr), idx);
#pragma unroll
                        for (int rr = 0; rr < FUSED_RB; rr++) acc[rr] += IT::dot(w[rr], idx & 1, acts);
                    } else {
                        // heavy items (27 words + realignment): one row at a time keeps the registers in bounds
#pragma unroll 1
                        for (int rr = 0; rr < FUSED_RB; rr++) {
                            typename IT::wts w;
                            IT::load(w, slot + row_off(rr), idx);
                            const float v = IT::dot(w, idx & 1, acts);
                            if (rr == 0) acc[0] += v; else acc[1] += v;
                        }
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
                    if (W == 1) batch_store_row(dst + lb * R + r, v, a, scatter);
                    else        pb[r * W + wg] = v;
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_a(empty + 8 * s);      // this warp is done with the stage
        if (W > 1 && g < G && c_first < nchunks) {
            // the W warps of the group meet; warp 0 of the group adds the partials in fixed order.  pb
            // alternates per working stage of the group: the buffer written two working stages later is
            // only touched after the barrier in between, which warp 0 reaches after these reads.
            { BATCH_ACC_T0(t_w);
            named_bar_sync(1 + g, W * 32);
            BATCH_ACC_ADD(acc_bar, t_w); }
            if (wg == 0) {
                for (int j = lane; ; j += 32) {
                    const int c = c_first + (j / FUSED_RB) * G;
                    const int r = c * FUSED_RB + (j % FUSED_RB);
                    if (c >= nchunks) break;
                    if (r < rows_here) {
                        float v = 0.0f;
                        for (int u = 0; u < W; u++) v += pb[r * W + u];
                        batch_store_row(dst + lb * R + r, v, a, scatter);
                    }
                }
            }
            git++;
        }
    }
}

// ---------------------------------------------------------------- the persistent kernel
// 16 warps = 4 per SM sub-partition: 4 warps share 16 K registers -> 128 per thread is the ceiling.
// One instantiation per format family: FAM 0 = Q4_K / Q5_K (16-byte aligned 128-weight items), FAM 1 = Q6_K,
// FAM 2 = Q4_0 / Q8_0 (four 32-weight blocks per item).
// Mixed in one kernel, Q6_K's 27-word realigned items push the loop state of ALL formats into local memory
// (measured: Q4_K falls from 5.8 to 2.8 TB/s); the nodes of a list are independent, so b200_mul_mat_batch
// simply sorts them into one launch per family.
template <int FAM>
__global__ void __launch_bounds__(BATCH_THREADS, 1)
k_mmvq_batch(const __grid_constant__ batch_args a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t * full       = reinterpret_cast<uint64_t *>(smem);
    uint64_t * empty      = full + BATCH_MAX_STAGES;
    uint64_t * acts_ready = full + 2 * BATCH_MAX_STAGES;
    uint64_t * acts_free  = acts_ready + BATCH_MAX_ABUF;
    unsigned * sdone      = reinterpret_cast<unsigned *>(acts_free + BATCH_MAX_ABUF);   // [8] consumer warps done with node i % 8
    float    * part       = reinterpret_cast<float *>(smem + a.off_part);
    uint8_t  * actbufs    = smem + a.off_acts;
    uint8_t  * slots      = smem + a.off_slots;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int S = a.S, NB = a.NB;

    if (tid == 0) {
        for (int s = 0; s < S; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], a.nwc); }
        for (int u = 0; u < NB; u++) { mbar_init(&acts_ready[u], 1); mbar_init(&acts_free[u], a.nwc); }
        for (int u = 0; u < 8; u++) sdone[u] = 0u;
        mbar_fence_init();
    }
    __syncthreads();
    pdl_launch();
    if (tid == 0) BATCH_STAMP(0, 6);

    if (warp == 0) {
        // ===================================================== weight producer: slabs of every node -> stage ring
        if (!a.const_weights) pdl_wait();
        int s = 0;
        uint32_t ph = 1;                                          // parity of the phase before the one to wait for; first lap: no wait
        bool lap0 = true;
        int slab_v = (int) blockIdx.x;
        node_cache cache;
        BATCH_ACC_DECL(acc_empty); BATCH_ACC_T0(t_begin);
        for (int i = 0; i < a.count; i++) {
            const node_r nd = cache.get(a, i, lane);
            int r_begin, nrows;
            batch_slab(nd, slab_v, r_begin, nrows);
            const int R = nd.R;
            const int nbatch = (nrows + R - 1) / R;
            const int row_pitch = batch_row_pitch(nd.row_bytes);
            for (int lb = 0; lb < nbatch; lb++) {
                { BATCH_ACC_T0(t_w);
                if (!lap0) mbar_wait(&empty[s], ph);              // the whole warp polls
                BATCH_ACC_ADD(acc_empty, t_w); }
                if (lane == 0) {
                    const int rows_here = (nrows - lb * R < R) ? (nrows - lb * R) : R;
                    const uint8_t * src = nd.w + (int64_t) (r_begin + lb * R) * nd.row_stride;
                    uint8_t * slot = slots + (size_t) s * a.slot_stride;
                    if (nd.contiguous) {
                        const uint32_t off = (uint32_t) (reinterpret_cast<uintptr_t>(src) & 15);
                        const uint32_t bytes = (off + (uint32_t) rows_here * (uint32_t) nd.row_bytes + 15u) & ~15u;
                        mbar_arrive_expect_tx(&full[s], bytes);
                        bulk_g2s(slot, src - off, bytes, &full[s]);
                    } else {
                        uint32_t total = 0;
                        for (int r = 0; r < rows_here; r++) {
                            const uint32_t off = (uint32_t) (reinterpret_cast<uintptr_t>(src + r * nd.row_stride) & 15);
                            total += (off + (uint32_t) nd.row_bytes + 15u) & ~15u;
                        }
                        mbar_arrive_expect_tx(&full[s], total);
                        for (int r = 0; r < rows_here; r++) {
                            const uint8_t * p = src + r * nd.row_stride;
                            const uint32_t off = (uint32_t) (reinterpret_cast<uintptr_t>(p) & 15);
                            bulk_g2s(slot + r * row_pitch, p - off, (off + (uint32_t) nd.row_bytes + 15u) & ~15u, &full[s]);
                        }
                    }
                }
                __syncwarp();
                if (++s == S) { s = 0; ph ^= 1u; lap0 = false; }
            }
        }
        if (lane == 0) { BATCH_STAMP(0, 7); BATCH_ACC_DECL(acc_total); BATCH_ACC_ADD(acc_total, t_begin); BATCH_ACC_OUT(95, 0, acc_empty); BATCH_ACC_OUT(95, 1, acc_total); }
        return;
    }

    if (warp == 1) {
        // ===================================================== activation producer: q8 activations of node i -> buffer i % NB
        pdl_wait();                                               // k_quantize_batch (and whatever made x) is done
        // arrivals a complete node has had: one per CTA; for a scattered node one per rank
        // (one-way scatter, epoch == NULL: the other device only receives rows, it takes no part in the counters)
        // (row split: the CTAs of a rank meet on a LOCAL arrival counter and the last one signals every rank once -- a counter
        //  that took an add from every CTA of every rank, 8 x 148 remote atomics on one word per node, cost ~20 us per link)
        const unsigned per_launch = a.epoch ? (unsigned) (a.n_peers + 1) : gridDim.x;
        unsigned * const done_set = a.done + (a.epoch ? (ld_relaxed_sys(a.epoch) & 1u) * BATCH_MAX_NODES : 0u);
        node_cache cache;
        int slab_v = (int) blockIdx.x;
        for (int i = 0; i < a.count; i++) {
            const node_r nd = cache.get(a, i, lane);
            int r_begin, nrows;
            batch_slab(nd, slab_v, r_begin, nrows);
            const int u = i % NB;
            if (i >= NB) mbar_wait(&acts_free[u], (uint32_t) (((i / NB) - 1) & 1));
            if (nd.dep >= 0) {
                // the node this one reads from (or must not overtake) is complete on EVERY CTA: acquire its counter.
                // All CTAs are resident (one per SM), so the wait ends; a bound turns a logic error into a trap, not a hang.
                if (lane == 0) {
                    // a scattered node is complete when every CTA of every rank has arrived, a local one when every CTA has
                    const bool dep_all = (a.nodes[nd.dep].qflags & BATCH_Q_SCATTER) != 0;
                    batch_wait_counter(done_set + nd.dep, dep_all ? per_launch : gridDim.x, a.epoch != nullptr);
                    BATCH_STAMP(i, 2);
                }
                __syncwarp();
            }
            if (lane == 0) {
                uint8_t * buf = actbufs + (size_t) u * a.acts_stride;
                cnode * hd = reinterpret_cast<cnode *>(buf);
                hd->dst = nd.dst; hd->k = nd.k; hd->pitch = nd.contiguous ? nd.row_bytes : batch_row_pitch(nd.row_bytes);
                hd->r_begin = r_begin; hd->nrows = nrows; hd->type = nd.type; hd->R = nd.R; hd->W = nd.W; hd->G = nd.G;
                hd->skew0 = (int) (reinterpret_cast<uintptr_t>(nd.w + (int64_t) r_begin * nd.row_stride) & 15);
                hd->rs16 = (int) (nd.row_stride & 15); hd->contiguous = nd.contiguous;
                hd->qflags = nd.qflags; hd->x = nd.x; hd->signal = done_set + i;
                if (nd.qflags & BATCH_Q_LOAD) {
                    const uint32_t bytes = batch_acts_bytes(nd.k);    // k % 256 == 0: a multiple of 16
                    mbar_arrive_expect_tx(&acts_ready[u], bytes);     // release: the header is visible to whoever sees the phase complete
                    bulk_g2s(buf + BATCH_AHDR, nd.acts, bytes, &acts_ready[u]);
                } else {
                    mbar_arrive_cta(&acts_ready[u]);                    // the consumers quantize x themselves, or keep their registers
                }
            }
            __syncwarp();
        }
        // row split: the launch is over when the last scattered node is complete on every rank -- CTA 0 stays for it
        if (a.epoch && a.n_peers > 0 && a.last_scatter >= 0 && blockIdx.x == 0) {
            if (lane == 0) batch_wait_counter(done_set + a.last_scatter, per_launch, true);
            __syncwarp();
        }
        return;
    }

    // ========================================================= consumers
    const int cw = warp - 2;
    int st_s = 0, cc = 0, layout_w = 0, git = 0;
    uint32_t st_ph = 0;
    const uint32_t full_a = smem_u32(full), empty_a = smem_u32(empty), ready_a = smem_u32(acts_ready), free_a = smem_u32(acts_free);
    const int slot_stride = a.slot_stride, count = a.count, acts_stride = a.acts_stride, nwc = a.nwc;
    acts128 acts = {};                                    // survives from node to node (BATCH_Q_SAME)
    BATCH_ACC_DECL(acc_full); BATCH_ACC_DECL(acc_bar); BATCH_ACC_DECL(acc_ready); BATCH_ACC_DECL(acc_layout); BATCH_ACC_T0(t_begin);
    for (int i = 0, u = 0, uph = 0; i < count; i++) {
        uint8_t * buf = actbufs + u * acts_stride;
        { BATCH_ACC_T0(t_w);
        mbar_wait_a(ready_a + 8 * u, (uint32_t) uph);
        BATCH_ACC_ADD(acc_ready, t_w); }
        const cnode nd = *reinterpret_cast<const cnode *>(buf);
        if (cw == 0 && lane == 0) BATCH_STAMP(i, 0);
        if (nd.qflags & BATCH_Q_INK) {
            batch_quantize_ink(nd.x, nd.k, buf + BATCH_AHDR, cw, lane, nwc);
            named_bar_sync(14, nwc * 32);                 // ids 1 .. 7 belong to the row groups (W > 1 -> G <= 7), 15 to the layout change
        }
        if (cw == 0 && lane == 0) BATCH_STAMP(i, 3);
#ifdef B200_TUNING
        const int st_s0 = st_s;
        if (cw == 0 && nd.nrows > 0) { mbar_wait_a(full_a + 8 * st_s0, st_ph); if (lane == 0) BATCH_STAMP(i, 4); }     // (a second wait on a completed phase returns at once)
#endif
        // the partials area is laid out per W and only rows that need several warps (W > 1) use it: as long
        // as every such node of the list has the same W, the groups, their barriers and their buffer
        // parity (git) simply carry on across nodes; a different W needs every consumer warp to be past its
        // last reduction before the new layout is written
        if (nd.W > 1) {
            if (layout_w != 0 && layout_w != nd.W) { BATCH_ACC_T0(t_w); named_bar_sync(BATCH_NWC + 1, a.nwc * 32); git = 0; BATCH_ACC_ADD(acc_layout, t_w); }
            layout_w = nd.W;
        }
        if (FAM == 2) {
            if (nd.type == T_Q8_0) batch_consume<T_Q8_0>(a, nd, acts, cw, lane, st_s, st_ph, cc, git, buf + BATCH_AHDR, free_a + 8 * u, full_a, empty_a, part, slots, S, slot_stride BATCH_TUN_ARGS);
            else                   batch_consume<T_Q4_0>(a, nd, acts, cw, lane, st_s, st_ph, cc, git, buf + BATCH_AHDR, free_a + 8 * u, full_a, empty_a, part, slots, S, slot_stride BATCH_TUN_ARGS);
        }
        else if (FAM == 1) { acts128 a6; batch_consume<T_Q6_K>(a, nd, a6, cw, lane, st_s, st_ph, cc, git, buf + BATCH_AHDR, free_a + 8 * u, full_a, empty_a, part, slots, S, slot_stride BATCH_TUN_ARGS); }
        else if (nd.type == T_Q5_K) batch_consume<T_Q5_K>(a, nd, acts, cw, lane, st_s, st_ph, cc, git, buf + BATCH_AHDR, free_a + 8 * u, full_a, empty_a, part, slots, S, slot_stride BATCH_TUN_ARGS);
        else                   batch_consume<T_Q4_K>(a, nd, acts, cw, lane, st_s, st_ph, cc, git, buf + BATCH_AHDR, free_a + 8 * u, full_a, empty_a, part, slots, S, slot_stride BATCH_TUN_ARGS);
        if (cw == 0 && lane == 0) BATCH_STAMP(i, 5);
        if (nd.qflags & BATCH_Q_SIGNAL) {
            // this warp's rows of node i are stored; the last of the CTA's consumer warps to get here releases them
            // to the other CTAs (cta-scope fences chain the warps' stores to the gpu-scope release)
            __syncwarp();
            if (lane == 0) {
                asm volatile("fence.acq_rel.cta;" ::: "memory");
                const unsigned old = atomicAdd(&sdone[i & 7], 1u);
                if (old == (unsigned) nwc - 1u) {
                    asm volatile("fence.acq_rel.cta;" ::: "memory");      // the other warps' arrivals, hence their stores, are observed ..
                    sdone[i & 7] = 0u;
                    if ((nd.qflags & BATCH_Q_SCATTER) && a.epoch) {        // .. and released (cumulatively) to every rank:
                        // this CTA's rows are performed system-wide (one fence), it arrives on the rank's local counter of
                        // the node (two sets behind the signal word), and the LAST CTA of the rank adds one to the node's
                        // counter on every rank (fence + relaxed add = a release pattern on each).  A red.release.sys per
                        // rank from every CTA made each link cost 44 us on 8 GPUs (profiles/r03p_bench_8gpu.json).
                        asm volatile("fence.acq_rel.sys;" ::: "memory");
                        unsigned prior;
                        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;" : "=r"(prior) : "l"(nd.signal + 2 * BATCH_MAX_NODES), "r"(1u) : "memory");
                        if (prior == gridDim.x - 1u) {
                            asm volatile("fence.acq_rel.sys;" ::: "memory");
                            red_relaxed_sys_add(nd.signal, 1u);
#pragma unroll 1
                            for (int r = 0; r < a.n_peers; r++)
                                red_relaxed_sys_add(reinterpret_cast<unsigned *>(reinterpret_cast<char *>(nd.signal) + a.peer_delta[r]), 1u);
                        }
                    } else {
                        red_release_gpu_add(nd.signal, 1u);                // .. and released (cumulatively) to the other CTAs
                    }
                    BATCH_STAMP(i, 1);
                }
            }
        }
        if (++u == NB) { u = 0; uph ^= 1; }
    }
    if (lane == 0 && (cw == 0 || cw == 9)) {
        BATCH_ACC_DECL(acc_total); BATCH_ACC_ADD(acc_total, t_begin);
        const int row = cw == 0 ? 94 : 93;
        BATCH_ACC_OUT(row, 0, acc_full); BATCH_ACC_OUT(row, 1, acc_bar); BATCH_ACC_OUT(row, 2, acc_ready); BATCH_ACC_OUT(row, 3, acc_layout); BATCH_ACC_OUT(row, 4, acc_total);
    }
}

} // namespace b200

namespace b200 {

// ================================================================= host side
// Can this MUL_MAT ride in a persistent launch?  (n = 1, plain 2-D, a format with a 128-weight item.)
static bool batch_node_ok(int type, const uint8_t * w, int64_t row_stride, int64_t m, int64_t k, const float * x) {
    if (!type_is_quant(type)) return false;
    if (m < 1 || m > 0x7FFFFFFF || k < 1024 || k % 256 != 0) return false;
    if (ceil_div(k / 128, 32) > 8) return false;                 // k <= 32768: one item per lane, same (W, lane) map as the per-node kernel
    if (reinterpret_cast<uintptr_t>(x) % 16 != 0) return false;
    if ((type == T_Q4_K || type == T_Q5_K) ? !hitem<T_Q4_K>::ok(k, row_stride, reinterpret_cast<uintptr_t>(w))
                                           : !hitem<T_Q6_K>::ok(k, row_stride, reinterpret_cast<uintptr_t>(w))) return false;
    const int64_t row_bytes = k / type_qk(type) * type_bytes(type);
    if (row_stride < row_bytes) return false;
    return true;
}

static inline size_t batch_acts_record_bytes(int64_t k) { return ((size_t) (k + k / 32 * 4) + 127) & ~(size_t) 127; }

// where the next q8 activation record goes in the caller's workspace (shared by the families' builders)
struct ws_cursor {
    uint8_t * p = nullptr;
    size_t left = 0;
    void reset(uint8_t * w, size_t bytes) {
        p = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(w) + 127) & ~(uintptr_t) 127);
        const size_t skew = (size_t) (p - w);
        left = (w && bytes > skew) ? bytes - skew : 0;
    }
};

static inline int batch_family(int type) { return type == T_Q6_K ? 1 : ((type == T_Q4_0 || type == T_Q8_0) ? 2 : 0); }

constexpr size_t BATCH_DONE_BYTES = (BATCH_MAX_NODES * sizeof(unsigned) + 127) & ~(size_t) 127;   // completion counters of one launch

struct batch_builder {
    batch_args args;
    int max_slot = 0;
    bool all_const = true;
    int  fam = 0;                     // format family of this builder's nodes (kernel instantiation)
    std::vector<int> orig;            // index of each node in the caller's list (fallback replay)
    explicit batch_builder(int family) : fam(family) { reset(); }
    int64_t bytes_w14 = 0, bytes_w12 = 0;                       // weight bytes of nodes that gain / do not gain from 14 consumer warps
    int waited = -1;                  // the latest node some earlier node of this launch already waits for
    void reset() {
        args.count = 0; args.kmax = 0; args.done = nullptr; args.epoch = nullptr; args.n_peers = 0; args.last_scatter = -1;
        max_slot = 0; all_const = true; bytes_w14 = bytes_w12 = 0; orig.clear(); waited = -1;
    }
    // row split: the family's control block in the local copy of the symmetric region (epoch word, then the counters)
    // and the byte distances to the other ranks' copies
    void set_peers(unsigned * ctrl, const long long * delta, int n_peers) {
        args.epoch = ctrl; args.done = ctrl + 16; args.n_peers = n_peers;
        for (int r = 0; r < n_peers; r++) args.peer_delta[r] = delta[r];
    }
    bool full() const { return args.count >= BATCH_MAX_NODES; }
    bool empty() const { return args.count == 0; }

    // dep: index (in this builder) of the latest node this one must wait for (it reads or overwrites what that node
    // writes, or overwrites what it reads), -1: none; x_from_dep: x is written by a node of this builder.
    // false: the workspace cannot hold this node's activation record
    bool add(int type, const uint8_t * w, int64_t row_stride, int64_t m, int64_t k, const float * x, float * dst, bool const_w, int grid,
             ws_cursor & cur, int dep = -1, bool x_from_dep = false, int orig_index = -1, bool scatter = false) {
        static const int slot_target = env_int("GGML_B200_BATCH_SLOT_BYTES", 46080);
        if (!args.done) {                                       // first node since reset: the launch's completion counters
            if (BATCH_DONE_BYTES > cur.left) return false;
            args.done = reinterpret_cast<unsigned *>(cur.p); cur.p += BATCH_DONE_BYTES; cur.left -= BATCH_DONE_BYTES;
        }
        batch_node & nd = args.nodes[args.count];
        const batch_node * prev = args.count > 0 ? &args.nodes[args.count - 1] : nullptr;
        const bool same_x = prev && prev->x == x && prev->k == k;
        // a CTA finishes nodes in order and the activation producer publishes them in order: once a node has waited for
        // node d, every node <= d is complete for all later nodes too -- they need no wait of their own (a poll under
        // load queues behind the SM's outstanding weight copies: measured ~3 us, profiles/r02e_chain_stamps.txt)
        nd.dep = (short) (dep > waited ? dep : -1);
        if (dep > waited) waited = dep;
        if (same_x && prev->type == type && fam != 1) {           // (Q6_K's 27-word items leave no registers to keep them in)
            // same activations and format as the previous node (q/k/v, gate/up): the consumers keep their registers.
            // (the previous node already waited for whatever wrote x; a hazard of this node's own dst still sets dep)
            nd.acts = prev->acts; nd.qflags = BATCH_Q_SAME;
        } else if (x_from_dep) {
            nd.acts = nullptr; nd.qflags = BATCH_Q_INK;         // quantized inside the launch, after node dep
        } else if (same_x && !(prev->qflags & BATCH_Q_INK) && prev->acts) {
            nd.acts = prev->acts; nd.qflags = BATCH_Q_LOAD;     // another format on the same activations: the record is shared
        } else {
            const size_t need = batch_acts_record_bytes(k);
            if (need > cur.left) return false;
            nd.acts = cur.p; nd.qflags = BATCH_Q_PRE | BATCH_Q_LOAD; cur.p += need; cur.left -= need;
        }
        if (nd.dep >= 0) args.nodes[nd.dep].qflags |= BATCH_Q_SIGNAL;
        if (scatter) { nd.qflags |= BATCH_Q_SCATTER; args.last_scatter = args.count; }
        orig.push_back(orig_index);
        args.count++;
        nd.w = w; nd.x = x; nd.dst = dst; nd.row_stride = row_stride; nd.m = (int) m; nd.k = (int) k;
        nd.row_bytes = (int) (k / type_qk(type) * type_bytes(type));
        nd.type = (unsigned char) type;
        nd.W = (unsigned char) ceil_div(k / 128, 32);
        nd.G = 0;                                               // batch_launch: consumer warps of the launch / W
        (nd.W == 7 || nd.W == 2 ? bytes_w14 : bytes_w12) += (int64_t) m * nd.row_bytes;
        // rows per stage: the nearest number of row pairs to the target; mat-vecs whose rows need several
        // warps (W > 1) keep their partials in shared memory, sized for BATCH_RMAX rows
        int R = 2 * ((slot_target + nd.row_bytes) / (2 * nd.row_bytes));
        if (R < 2) R = 2;
        if (R > (nd.W > 1 ? BATCH_RMAX : BATCH_RMAX_W1)) R = nd.W > 1 ? BATCH_RMAX : BATCH_RMAX_W1;
        nd.R = (unsigned char) R;
        nd.contiguous = row_stride == nd.row_bytes ? 1 : 0;
        int rpc = (int) ceil_div(m, grid);
        rpc += rpc & 1;
        nd.rows_per_cta = rpc;
        // equal stages: a 28-row slab as 14 + 14 rows instead of 20 + 8 (the short stage costs as much fixed work
        // and leaves most consumer warps without a row pair)
        {
            const int ns = (rpc + R - 1) / R;
            int Re = (rpc + ns - 1) / ns;
            Re += Re & 1;
            if (Re < R) R = Re;
            nd.R = (unsigned char) R;
        }
        const int row_pitch = (nd.row_bytes + 15 + 15 + 16) & ~15;
        const int slot_bytes = nd.contiguous ? (15 + R * nd.row_bytes + 15 + 16) : R * row_pitch;
        if (slot_bytes > max_slot) max_slot = slot_bytes;
        if (nd.k > args.kmax) args.kmax = nd.k;
        all_const = all_const && const_w;
        return true;
    }
};

#ifdef B200_TUNING
static unsigned long long * g_batch_dbg = nullptr;   // set through b200_debug_batch_stamps (tuning builds only)
#endif
static int batch_launch(batch_builder & bb, cudaStream_t st) {
    if (bb.empty()) return B200_OK;
#ifdef B200_TUNING
    bb.args.dbg = g_batch_dbg;
#endif
    const dev_caps & caps = current_caps();
    batch_args & a = bb.args;
    static const int s_env = env_int("GGML_B200_BATCH_STAGES", BATCH_MAX_STAGES);
    static const int nb_env = env_int("GGML_B200_BATCH_ABUF", 2);
    static const int pdl = env_int("GGML_B200_FUSED_PDL", 1);
    a.const_weights = bb.all_const ? 1 : 0;
    if (a.last_scatter >= 0) a.nodes[a.last_scatter].qflags |= BATCH_Q_SIGNAL;      // the closing barrier of a row-split launch
    // consumer warps: 12 (14 warps = 448 threads) is 1-1.5 % faster on one-warp rows (k <= 4096) than 14; rows of
    // 7 warps (k = 28672, the Llama-3-70B down projection) leave 5 of 12 idle but fill 14 (60 % -> 85 % of the
    // measured HBM peak), rows of 2 warps gain 2 %.  Both are within the 128-register ceiling (<= 4 warps / SMSP).
    static const int nwc_env = env_int("GGML_B200_BATCH_NWC", 0);
    // A launch with chain links is bound by the LATENCY of a node on one CTA (a dependent node cannot start before the
    // previous one has finished everywhere): 14 warps take the 14 row pairs of a 4096-row slab in one round instead of two
    // (measured, profiles/r02c_chain_stamps.txt: 1.25 us per round of row pairs).
    static const int chain_nwc = env_int("GGML_B200_CHAIN_NWC", 14);
    bool chained = false;
    for (int i = 0; i < a.count; i++) chained = chained || a.nodes[i].dep >= 0;
    a.nwc = nwc_env == 12 || nwc_env == 14 ? nwc_env : (chained ? chain_nwc : (bb.bytes_w14 > bb.bytes_w12 ? 14 : 12));
    for (int i = 0; i < a.count; i++) {
        const int G = a.nwc / a.nodes[i].W;
        a.nodes[i].G = (unsigned char) (a.nodes[i].W == 1 && G > 12 && !chained ? 12 : G);
    }
    a.slot_stride = (bb.max_slot + 127) & ~127;
    a.off_part    = (2 * BATCH_MAX_STAGES + 2 * BATCH_MAX_ABUF) * 8 + 8 * (int) sizeof(unsigned);
    a.off_acts    = (a.off_part + 2 * BATCH_NWC * BATCH_RMAX * (int) sizeof(float) + 127) & ~127;
    a.acts_stride = BATCH_AHDR + (int) batch_acts_record_bytes(a.kmax);
    int NB = nb_env < 2 ? 2 : (nb_env > BATCH_MAX_ABUF ? BATCH_MAX_ABUF : nb_env);
    while (NB > 2 && NB * a.acts_stride > 49152) NB--;           // keep the weight ring large
    a.NB = NB;
    a.off_slots   = a.off_acts + NB * a.acts_stride;
    int S = (caps.smem_optin - a.off_slots) / a.slot_stride;
    if (S > s_env) S = s_env;
    if (S > BATCH_MAX_STAGES) S = BATCH_MAX_STAGES;
    if (S < 2) return B200_STREAM_NOT_APPLICABLE;
    a.S = S;
    const size_t smem = (size_t) a.off_slots + (size_t) S * a.slot_stride;

    static smem_limit lim[3];
    auto kern = bb.fam == 2 ? k_mmvq_batch<2> : (bb.fam == 1 ? k_mmvq_batch<1> : k_mmvq_batch<0>);
    if (const int rc = lim[bb.fam].ensure(kern, smem)) return rc;
    // 1. every activation vector of the list -> q8 (one launch)
    {
        const dim3 grid((unsigned) ceil_div(a.kmax / 128, 8), (unsigned) a.count);
        k_quantize_batch<<<grid, 256, 0, st>>>(a);
        b200_count_launch(1);
    }
    // 2. the persistent mat-vec; starts (and prefetches constant weights) while the quantizer drains
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) caps.sm_count); cfg.blockDim = dim3((unsigned) (2 + a.nwc) * 32);
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

} // namespace b200
