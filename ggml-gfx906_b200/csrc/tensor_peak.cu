// tensor_peak.cu — a tcgen05 micro-kernel that measures what the 5th-gen tensor cores of THIS box deliver:
// kind::i8 (int8 x int8 -> s32, the rate BASELINE.json's "batched MMQ int8 TOPS" term is a fraction of) and kind::f16
// (bf16 x bf16 -> f32, cross-check against the driver's cuBLAS figure in MEASURED_PEAKS.json).
//
// One CTA per SM, operands resident in shared memory (one 128 x 128-byte A tile, one 256 x 128-byte B tile, K-major,
// SWIZZLE_128B), one thread issuing `iters` x 4 MMAs (M = 128, N = 256, K = 32 bytes each) into one TMEM accumulator
// with no loads in between: nothing but the tensor pipe limits it.  The same kernel with iters = 1 writes the
// accumulator out, so a test can check the arithmetic (and with it the instruction descriptor) against the host.
// The reference has no counterpart (it never measures a peak); SURVEY §8(d) asks for this number.
#include "common.cuh"
#include "mmq_tc.cuh"

namespace b200 {

constexpr int PEAK_M = 128, PEAK_N = 256, PEAK_KB = 128;      // K extent of the resident tiles in BYTES (one swizzle row)

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
        "}\n" :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}

// a, b: row-major [rows][128] bytes (int8, or 64 bf16); out: [128][256] 32-bit accumulators (NULL: timing only)
template <bool I8>
__global__ void __launch_bounds__(128, 1) k_tensor_peak(const uint8_t * __restrict__ a, const uint8_t * __restrict__ b, uint32_t * __restrict__ out, int iters) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    uint8_t * sa = smem, * sb = smem + PEAK_M * PEAK_KB;
    uint64_t * bar = reinterpret_cast<uint64_t *>(sb + PEAK_N * PEAK_KB);
    uint32_t * tmem_slot = reinterpret_cast<uint32_t *>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5;

    // operands -> the UMMA canonical K-major SWIZZLE_128B layout: 16-byte chunk c of row r sits at chunk c ^ (r & 7)
    for (int i = tid; i < (PEAK_M + PEAK_N) * 8; i += 128) {
        const int r = i >> 3, c = i & 7;
        const bool is_a = r < PEAK_M;
        const int rr = is_a ? r : r - PEAK_M;
        const uint4 v = *reinterpret_cast<const uint4 *>((is_a ? a : b) + (size_t) rr * PEAK_KB + 16 * c);
        *reinterpret_cast<uint4 *>((is_a ? sa : sb) + rr * PEAK_KB + ((c ^ (rr & 7)) << 4)) = v;
    }
    if (tid == 0) { mbar_init(bar, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(tmem_slot, PEAK_N);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (tid == 0) {
        // D format: s32 (2) / f32 (1); A, B format: s8 (1) / bf16 (1); both K-major; N >> 3; M >> 4
        constexpr uint32_t idesc = ((I8 ? 2u : 1u) << 4) | (1u << 7) | (1u << 10) | ((uint32_t) (PEAK_N >> 3) << 17) | ((uint32_t) (PEAK_M >> 4) << 24);
        const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sb);
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int k4 = 0; k4 < 4; k4++) {            // 32 bytes (2 x 16-byte units) of K per MMA: 32 int8 or 16 bf16
                if (I8) umma_i8(tmem, da + (uint64_t) (2 * k4), db + (uint64_t) (2 * k4), idesc, (it > 0 || k4 > 0) ? 1u : 0u);
                else    umma_f16(tmem, da + (uint64_t) (2 * k4), db + (uint64_t) (2 * k4), idesc, (it > 0 || k4 > 0) ? 1u : 0u);
            }
        }
        umma_commit(bar);
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    if (out && blockIdx.x == 0) {
        // warp w reads TMEM lanes 32 w .. 32 w + 31 = accumulator rows
        const uint32_t taddr = tmem + ((uint32_t) (warp * 32) << 16);
        for (int c0 = 0; c0 < PEAK_N; c0 += 8) {
            uint32_t v[8];
            tmem_ld8(taddr + (uint32_t) c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 8; j++) out[(size_t) tid * PEAK_N + c0 + j] = v[j];
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, PEAK_N); }
}

} // namespace b200

using namespace b200;

// kind: 0 = int8 (s8 x s8 -> s32), 1 = bf16 (-> f32).  a_dev / b_dev: [128][128] and [256][128] bytes.
// out_dev != NULL: iters is forced to 1 and CTA 0 writes its [128][256] accumulators (verification).
// ops_per_s (may be NULL): 2 x MACs per second over all SMs, timed with CUDA events over `reps` launches.
extern "C" int b200_tensor_peak(int kind, const void * a_dev, const void * b_dev, void * out_dev, int iters, int reps,
                                double * ops_per_s, void * stream) {
    if (kind != 0 && kind != 1) return B200_ERR_TYPE;
    if (!a_dev || !b_dev || iters < 1 || reps < 1) return B200_ERR_SHAPE;
    cudaStream_t st = (cudaStream_t) stream;
    const dev_caps & caps = current_caps();
    const size_t smem = (size_t) (PEAK_M + PEAK_N) * PEAK_KB + 64 + 1024;
    auto kern = kind == 0 ? k_tensor_peak<true> : k_tensor_peak<false>;
    static smem_limit lim[2];
    if (const int rc = lim[kind].ensure(kern, smem)) return rc;
    if (out_dev) iters = 1;
    const int grid = out_dev ? 1 : caps.sm_count;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<<<grid, 128, smem, st>>>((const uint8_t *) a_dev, (const uint8_t *) b_dev, (uint32_t *) out_dev, iters);   // warm-up
    cudaEventRecord(e0, st);
    for (int r = 0; r < reps; r++) kern<<<grid, 128, smem, st>>>((const uint8_t *) a_dev, (const uint8_t *) b_dev, (uint32_t *) out_dev, iters);
    cudaEventRecord(e1, st);
    b200_count_launch((uint64_t) reps + 1);
    cudaError_t e = cudaEventSynchronize(e1);
    float ms = 0.0f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (e != cudaSuccess) return -(int) e;
    if (ops_per_s) {
        const double k_per_mma = kind == 0 ? 32.0 : 16.0;
        const double ops = 2.0 * PEAK_M * PEAK_N * k_per_mma * 4.0 * (double) iters * (double) grid * (double) reps;
        *ops_per_s = ms > 0.0f ? ops / (ms * 1e-3) : 0.0;
    }
    return B200_OK;
}
