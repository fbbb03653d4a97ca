// no_alloc.cu — a decode MUL_MAT entry point for callers that cannot allocate device memory (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  It has no counterpart in the
// reference.
//
// b200_mul_mat_vec_static is b200_quantize_acts + b200_mul_mat_vec_q with the activation scratch (int8 plane + scale
// plane, b200_acts_scratch_bytes(k, n) bytes) taken from a file-scope __device__ array instead of a caller buffer:
//   * the array is B200_STATIC_SCRATCH_BYTES (1 MiB) on every GPU that loads the library, until the process exits;
//   * problems that need more return B200_ERR_TOO_LARGE; nothing in this file calls cudaMalloc / cudaFree or any
//     allocating routine;
//   * every call on a device shares the array.  It keeps the previous call's contents; nothing has to start at zero
//     (the quantizer writes every byte the mat-vec reads: n * k int8 and n * k / 32 scales), but two calls whose GPU
//     work overlaps (different streams, or two host threads on one device) race on it — serialize them.
// In host code the array's NAME is a host-side stand-in, not a device address, so the address is fetched once per
// device with cudaGetSymbolAddress (that device current) and only that pointer is handed to kernels.
#include "common.cuh"
#include "mmvq_stream.cuh"

#include <atomic>

namespace b200 {

static __device__ __align__(256) unsigned char g_static_scratch[B200_STATIC_SCRATCH_BYTES];

// device address of g_static_scratch on the current device; NULL + *rc on failure
static unsigned char * static_scratch_ptr(int * rc) {
    static unsigned char * ptr[64];
    static std::atomic<int> have[64];
    const int dev = current_device_slot();
    if (have[dev].load(std::memory_order_acquire) == 2) return ptr[dev];
    void * p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) { *rc = e != cudaSuccess ? -(int) e : B200_ERR_WORKSPACE; return nullptr; }
    int expect = 0;
    if (have[dev].compare_exchange_strong(expect, 1, std::memory_order_acq_rel)) {
        ptr[dev] = (unsigned char *) p;
        have[dev].store(2, std::memory_order_release);
    }
    return (unsigned char *) p;                               // the same address whichever thread published it
}

} // namespace b200

using namespace b200;

extern "C" size_t b200_static_scratch_bytes(void) { return B200_STATIC_SCRATCH_BYTES; }

extern "C" int b200_mul_mat_vec_static(int type, const void * w_dev, size_t w_row_stride_bytes, int64_t m, int64_t k,
                                       const float * x_dev, size_t x_row_stride_bytes, int64_t n,
                                       float * dst_dev, size_t dst_col_stride_bytes, void * stream) {
    if (!type_is_quant(type)) return B200_ERR_TYPE;
    if (m < 0 || k < 0 || n < 0 || n > B200_MMVQ_MAX_N || k % type_qk(type) != 0) return B200_ERR_SHAPE;
    if (dst_col_stride_bytes % 4 != 0 || w_row_stride_bytes % 2 != 0 || x_row_stride_bytes % 4 != 0) return B200_ERR_STRIDE;
    if (b200_acts_scratch_bytes(k, n) > (size_t) B200_STATIC_SCRATCH_BYTES) return B200_ERR_TOO_LARGE;
    if (m == 0 || n == 0) return B200_OK;
    int rc = B200_OK;
    unsigned char * scratch = static_scratch_ptr(&rc);
    if (!scratch) return rc;
    if (k > 0) {
        rc = b200_quantize_acts(x_dev, x_row_stride_bytes, k, n, scratch, stream);
        if (rc) return rc;
    }
    return b200_mul_mat_vec_q(type, w_dev, w_row_stride_bytes, m, k, scratch, n, dst_dev, dst_col_stride_bytes, stream);
}
