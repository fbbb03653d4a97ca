// capi.cu — the non-kernel part of the C ABI (include/ggml_b200.h): version, error strings, device
// enumeration, the format table and the launch counter.
#include "common.cuh"

#include <atomic>
#include <cstdio>
#include <cstring>

using namespace b200;

static std::atomic<uint64_t> g_launches{0};

extern "C" void b200_count_launch(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" uint64_t b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int b200_abi_version(void) { return 1; }

extern "C" const char * b200_error_string(int code) {
    if (code < 0) return cudaGetErrorString((cudaError_t) (-code));
    switch (code) {
        case B200_OK:            return "ok";
        case B200_ERR_TYPE:      return "unsupported tensor type";
        case B200_ERR_SHAPE:     return "bad shape";
        case B200_ERR_STRIDE:    return "bad stride / alignment";
        case B200_ERR_WORKSPACE: return "workspace missing or too small";
        case B200_ERR_NO_DEVICE: return "no sm_100 device";
        default:                 return "unknown error";
    }
}

// Devices this library can run on: compute capability 10.x (the kernels are built for sm_100a only).
extern "C" int b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    int ok = 0;
    for (int i = 0; i < n; i++) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, i) == cudaSuccess && major == 10) ok++;
        else break;   // device ids stay dense: stop at the first non-Blackwell device
    }
    return ok;
}

extern "C" int b200_device_info(int device, char * name, size_t name_len, size_t * free_bytes, size_t * total_bytes,
                                int * sm_count, int * cc) {
    if (device < 0 || device >= b200_device_count()) return B200_ERR_NO_DEVICE;
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return -(int) e;
    if (name && name_len) { strncpy(name, prop.name, name_len - 1); name[name_len - 1] = 0; }
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (cc) *cc = prop.major * 100 + prop.minor * 10;
    if (free_bytes || total_bytes) {
        int prev = 0;
        cudaGetDevice(&prev);
        cudaSetDevice(device);
        size_t f = 0, t = 0;
        e = cudaMemGetInfo(&f, &t);
        cudaSetDevice(prev);
        if (e != cudaSuccess) return -(int) e;
        if (free_bytes) *free_bytes = f;
        if (total_bytes) *total_bytes = t;
    }
    return B200_OK;
}

extern "C" int    b200_block_elems(int type) { return type_qk(type); }
extern "C" size_t b200_block_bytes(int type) { return (size_t) type_bytes(type); }
extern "C" size_t b200_row_bytes(int type, int64_t k) {
    const int qk = type_qk(type);
    if (qk == 0 || k < 0 || k % qk != 0) return 0;
    return (size_t) (k / qk) * (size_t) type_bytes(type);
}
