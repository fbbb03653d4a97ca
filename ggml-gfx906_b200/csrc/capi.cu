// capi.cu — the non-kernel part of the C ABI (include/ggml_b200.h): version, error strings, device
// enumeration, the format table and the launch counter.
#include "common.cuh"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

using namespace b200;

static std::atomic<uint64_t> g_launches{0};

extern "C" void b200_count_launch(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" uint64_t b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int b200_abi_version(void) { return 2; }

// ---------------------------------------------------------------- run-time options
// Path switches the tests and the benches flip (one kernel against another on the same inputs).  Each starts
// from its environment variable, read ONCE; after that b200_set_option is the only way to change it, so no
// launch path ever calls getenv.
namespace {
struct opt_def { const char * env; int dflt; };
const opt_def g_opt_defs[B200_OPT_COUNT] = {
    /* B200_OPT_MMVQ_GENERIC        */ { "GGML_B200_MMVQ_GENERIC", 0 },
    /* B200_OPT_MMVQ_NO_FUSED       */ { "GGML_B200_MMVQ_NO_FUSED", 0 },
    /* B200_OPT_BATCH_SINGLE        */ { "GGML_B200_BATCH_SINGLE", -1 },
    /* B200_OPT_NO_BATCH_KERNEL     */ { "GGML_B200_NO_BATCH_KERNEL", 0 },
    /* B200_OPT_MMQ_BN              */ { "GGML_B200_MMQ_BN", 0 },
    /* B200_OPT_MMQ_FORCE_UNALIGNED */ { "GGML_B200_MMQ_FORCE_UNALIGNED", 0 },
    /* B200_OPT_NO_CHAIN            */ { "GGML_B200_NO_CHAIN", 0 },
    /* B200_OPT_MMQ_ROUTE           */ { "GGML_B200_MMQ_ROUTE", 0 },
};
std::atomic<int> g_opt[B200_OPT_COUNT];
std::once_flag   g_opt_once;
void opt_init() {
    std::call_once(g_opt_once, [] {
        for (int i = 0; i < B200_OPT_COUNT; i++) {
            const char * v = getenv(g_opt_defs[i].env);
            g_opt[i].store(v && *v ? atoi(v) : g_opt_defs[i].dflt, std::memory_order_relaxed);
        }
    });
}
}
extern "C" int b200_option(int id) {
    if (id < 0 || id >= B200_OPT_COUNT) return 0;
    opt_init();
    return g_opt[id].load(std::memory_order_relaxed);
}
extern "C" int b200_set_option(int id, int value) {
    if (id < 0 || id >= B200_OPT_COUNT) return 0;
    opt_init();
    return g_opt[id].exchange(value, std::memory_order_relaxed);
}

// ---------------------------------------------------------------- sticky device-side error flags
// Kernels cannot return codes: the ones that can meet bad DATA (a GET_ROWS index outside the table) set a bit in a
// word of mapped pinned host memory; b200_device_error_flags() reads it on the host (meaningful once the stream
// that ran the kernel has been synchronized).
namespace {
std::mutex g_err_mutex;
unsigned * g_err_host = nullptr;
}
extern "C" unsigned * b200_error_word_dev(void) {
    std::lock_guard<std::mutex> lock(g_err_mutex);
    if (!g_err_host) {
        void * p = nullptr;
        if (cudaHostAlloc(&p, 64, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        memset(p, 0, 64);
        g_err_host = (unsigned *) p;
    }
    void * d = nullptr;
    if (cudaHostGetDevicePointer(&d, g_err_host, 0) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return (unsigned *) d;
}
extern "C" unsigned b200_device_error_flags(int clear) {
    std::lock_guard<std::mutex> lock(g_err_mutex);
    if (!g_err_host) return 0;
    const unsigned v = *(volatile unsigned *) g_err_host;
    if (clear) *(volatile unsigned *) g_err_host = 0;
    return v;
}

extern "C" const char * b200_error_string(int code) {
    if (code < 0) return cudaGetErrorString((cudaError_t) (-code));
    switch (code) {
        case B200_OK:            return "ok";
        case B200_ERR_TYPE:      return "unsupported tensor type";
        case B200_ERR_SHAPE:     return "bad shape";
        case B200_ERR_STRIDE:    return "bad stride / alignment";
        case B200_ERR_WORKSPACE: return "workspace missing or too small";
        case B200_ERR_NO_DEVICE: return "no sm_100 device";
        case B200_ERR_INDEX:     return "row index outside the table";
        case B200_ERR_TOO_LARGE: return "problem exceeds the static scratch (B200_STATIC_SCRATCH_BYTES): use the entry point that takes a caller workspace";
        case B200_ERR_FILE:      return "GGUF file could not be opened, parsed or read (see b200_gguf_last_error)";
        case B200_NOT_APPLICABLE: return "no kernel of this library applies to the call";
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
