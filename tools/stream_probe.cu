// stream_probe.cu — tuning aid, not product code: how fast can one B200 stream a large buffer from HBM
// into the SMs with each copy mechanism?  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_probe stream_probe.cu -lcuda
//   mode 0: cp.async.bulk (1-D TMA), one issuing thread per CTA, ring of D slots of X bytes
//   mode 1: cp.async.bulk, D issuing threads (one per slot, different warps)
//   mode 2: cp.async 16 B (LDGSTS) by all threads into a ring, wait_group
//   mode 3: LDG.128 into registers, U loads in flight per thread
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t * bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect(uint64_t * bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t * bar, uint32_t parity) {
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n"
                 :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void * dst, const void * src, uint32_t bytes, uint64_t * bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- modes 0/1: bulk copies.  Each CTA streams a contiguous slab of `per_cta` bytes.
__global__ void k_bulk(const uint8_t * src, size_t per_cta, int X, int D, int multi, unsigned * sink) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t * full = reinterpret_cast<uint64_t *>(smem);
    uint8_t * slots = smem + 256;
    const uint8_t * base = src + (size_t) blockIdx.x * per_cta;
    const int n = (int) (per_cta / X);
    if (threadIdx.x == 0) {
        for (int s = 0; s < D; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned acc = 0;
    if (!multi) {
        // thread 0 issues and re-issues; all threads wait on each slot (consumers do nothing)
        if (threadIdx.x == 0) for (int b = 0; b < D && b < n; b++) { mbar_expect(&full[b], X); bulk_g2s(slots + (size_t) b * X, base + (size_t) b * X, X, &full[b]); }
        for (int b = 0; b < n; b++) {
            const int s = b % D;
            mbar_wait(&full[s], (b / D) & 1);
            acc += slots[(size_t) s * X + threadIdx.x];
            __syncthreads();
            if (threadIdx.x == 0 && b + D < n) { mbar_expect(&full[s], X); bulk_g2s(slots + (size_t) s * X, base + (size_t) (b + D) * X, X, &full[s]); }
        }
    } else {
        // warp w owns slots w, w + nw, ...: its lane 0 issues, the warp waits, lane 0 re-issues
        const int nw = blockDim.x >> 5;
        for (int b = warp; b < n; b += nw) {
            const int s = b % D;      // D == nw here
            if (lane == 0) { mbar_expect(&full[s], X); bulk_g2s(slots + (size_t) s * X, base + (size_t) b * X, X, &full[s]); }
            mbar_wait(&full[s], (b / D) & 1);
            acc += slots[(size_t) s * X + lane];
            __syncwarp();
        }
    }
    if (acc == 0xFFFFFFFFu) *sink = acc;
}

// ---- mode 4: the fused kernel's protocol: producer warp + full/empty mbarriers, NWC consumer warps that only
// wait, touch one byte and arrive.  poll_all: every lane polls (else lane 0 + __syncwarp).
__device__ __forceinline__ void mbar_arrive(uint64_t * bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__global__ void k_ring(const uint8_t * src, size_t per_cta, int X, int D, int NWC, int poll_all, unsigned * sink) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t * full = reinterpret_cast<uint64_t *>(smem);
    uint64_t * empty = full + 16;
    uint8_t * slots = smem + 256;
    const uint8_t * base = src + (size_t) blockIdx.x * per_cta;
    const int n = (int) (per_cta / X);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < D; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], NWC); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp == NWC) {
        if (lane == 0) for (int b = 0; b < n; b++) {
            const int s = b % D;
            if (b >= D) mbar_wait(&empty[s], ((b / D) - 1) & 1);
            mbar_expect(&full[s], X); bulk_g2s(slots + (size_t) s * X, base + (size_t) b * X, X, &full[s]);
        }
        return;
    }
    unsigned acc = 0;
    for (int b = 0; b < n; b++) {
        const int s = b % D;
        if (poll_all || lane == 0) mbar_wait(&full[s], (b / D) & 1);
        __syncwarp();
        acc += slots[(size_t) s * X + threadIdx.x];
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
    if (acc == 0xFFFFFFFFu) *sink = acc;
}

// ---- mode 2: cp.async 16 B by every thread, ring of D groups of X bytes
template <int D>
__global__ void k_ldgsts(const uint8_t * src, size_t per_cta, int X, unsigned * sink) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t * slots = smem;
    const uint8_t * base = src + (size_t) blockIdx.x * per_cta;
    const int n = (int) (per_cta / X);
    unsigned acc = 0;
    auto issue = [&](int b) {
        const uint8_t * g = base + (size_t) b * X;
        uint8_t * s = slots + (size_t) (b % D) * X;
        for (int o = threadIdx.x * 16; o < X; o += blockDim.x * 16)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_u32(s + o)), "l"(g + o) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int b = 0; b < D - 1; b++) { if (b < n) issue(b); else asm volatile("cp.async.commit_group;" ::: "memory"); }
    for (int b = 0; b < n; b++) {
        if (b + D - 1 < n) issue(b + D - 1); else asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group %0;" :: "n"(D - 1) : "memory");
        acc += slots[(size_t) (b % D) * X + threadIdx.x];
        __syncthreads();
    }
    if (acc == 0xFFFFFFFFu) *sink = acc;
}

// ---- mode 3: LDG.128, U independent loads per thread per iteration
template <int U>
__global__ void k_ldg(const uint4 * src, size_t per_cta16, unsigned * sink) {
    const uint4 * base = src + (size_t) blockIdx.x * per_cta16;
    unsigned acc = 0;
    for (size_t i = threadIdx.x; i + (size_t) (U - 1) * blockDim.x < per_cta16; i += (size_t) U * blockDim.x) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = __ldcs(base + i + (size_t) u * blockDim.x);
#pragma unroll
        for (int u = 0; u < U; u++) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678u) *sink = acc;
}

int main(int argc, char ** argv) {
    const size_t total = (size_t) 2 << 30;
    uint8_t * buf; unsigned * sink;
    cudaMalloc(&buf, total); cudaMalloc(&sink, 4);
    cudaMemset(buf, 1, total);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto report = [&](const char * what, float ms, size_t bytes) { printf("%-60s %8.1f GB/s\n", what, bytes / ms / 1e6); fflush(stdout); };
    char name[256];
    cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(k_ldgsts<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(k_ldgsts<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(k_ldgsts<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    for (int multi = 0; multi < 2; multi++)
    for (int X : { 4096, 8192, 16384, 32768, 65536 })
    for (int D : { 2, 4, 8, 16 }) {
        if ((size_t) X * D > 200 * 1024) continue;
        const int grid = sms;
        const size_t per_cta = total / grid / X * X;
        const int threads = multi ? D * 32 : 128;
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            k_bulk<<<grid, threads, 256 + (size_t) X * D>>>(buf, per_cta, X, D, multi, sink);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        snprintf(name, sizeof name, "bulk copy %s X=%6d D=%2d (1 CTA/SM)", multi ? "issuer/slot" : "1 issuer   ", X, D);
        report(name, ms, per_cta * grid);
        if (cudaGetLastError() != cudaSuccess) { printf("error\n"); return 1; }
    }
    cudaFuncSetAttribute(k_ring, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    for (int poll_all = 0; poll_all < 2; poll_all++)
    for (int X : { 4096, 4608, 8192, 9216, 18432 })
    for (int D : { 4, 8, 16 })
    for (int NWC : { 1, 4, 9 }) {
        if ((size_t) X * D > 200 * 1024) continue;
        const int grid = sms;
        const size_t per_cta = total / grid / X * X;
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            k_ring<<<grid, (NWC + 1) * 32, 256 + (size_t) X * D>>>(buf, per_cta, X, D, NWC, poll_all, sink);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        snprintf(name, sizeof name, "ring: producer warp + %d consumer warps, X=%6d D=%2d %s", NWC, X, D, poll_all ? "all lanes poll" : "lane 0 polls");
        report(name, ms, per_cta * grid);
        if (cudaGetLastError() != cudaSuccess) { printf("error\n"); return 1; }
    }
    for (int ctas : { 1, 2, 4 })
    for (int threads : { 256, 512, 1024 }) {
        if (ctas * threads > 2048) continue;
        const int grid = sms * ctas;
        const size_t per_cta16 = total / 16 / grid;
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            k_ldg<8><<<grid, threads>>>((const uint4 *) buf, per_cta16, sink);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        snprintf(name, sizeof name, "LDG.128 x8 in flight, %d CTA/SM x %d threads", ctas, threads);
        report(name, ms, per_cta16 * 16 * grid);
    }
    for (int X : { 8192, 16384, 32768 })
    for (int D : { 2, 4, 6 }) {
        if ((size_t) X * D > 200 * 1024) continue;
        const int grid = sms;
        const size_t per_cta = total / grid / X * X;
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            if (D == 2) k_ldgsts<2><<<grid, 256, (size_t) X * D>>>(buf, per_cta, X, sink);
            else if (D == 4) k_ldgsts<4><<<grid, 256, (size_t) X * D>>>(buf, per_cta, X, sink);
            else k_ldgsts<6><<<grid, 256, (size_t) X * D>>>(buf, per_cta, X, sink);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        snprintf(name, sizeof name, "cp.async 16B ring X=%6d D=%d, 256 thr", X, D);
        report(name, ms, per_cta * grid);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
