// plugin.cpp — libggml-b200.so: the ggml backend plugin around the B200 kernel library.
//
// An unmodified ggml loads this with ggml_backend_load() / GGML_BACKEND_PATH
// (reference src/ggml-backend-reg.cpp:239-283, 595-599).  It fills the reference's own vtables
// (src/ggml-backend-impl.h:17-207) and plays the role of src/ggml-cuda/ggml-cuda.cu for the ops it
// claims: MUL_MAT with Q4_0/Q8_0/Q4_K/Q5_K/Q6_K weights and f32 activations, GET_ROWS on those
// formats, CPY/DUP/CONT from those formats to f32/f16.  Everything else is answered with
// supports_op = false so ggml_backend_sched keeps it on another backend; nothing claimed here ever
// falls back to the CPU.
//
// Host-side design (not a port of ggml-cuda.cu): one stream per backend instance, one grow-only
// workspace per backend, every kernel reached through the C ABI of include/ggml_b200.h, and whole
// graph_compute() calls replayed as CUDA graphs (see graph_cache below) because decode graphs are
// hundreds of 1-5 us kernels.
#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-backend-impl.h"

#include "../../include/ggml_b200.h"
#include "../../include/ggml_b200_plugin.h"

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#define B200_MAX_DEVICES 16

static void b200_log(const char * fmt, ...) __attribute__((format(printf, 1, 2)));
static void b200_log(const char * fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    fprintf(stderr, "ggml-b200: ");
    vfprintf(stderr, fmt, ap);
    va_end(ap);
}

// Device errors are fatal, as in the reference (CUDA_CHECK -> GGML_ABORT, src/ggml-cuda/common.cuh:145-156)
#define CUDA_OK(expr)                                                                                  \
    do {                                                                                               \
        cudaError_t e_ = (expr);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            b200_log("%s failed: %s (%s:%d)\n", #expr, cudaGetErrorString(e_), __FILE__, __LINE__);    \
            GGML_ABORT("ggml-b200: CUDA error");                                                       \
        }                                                                                              \
    } while (0)
#define B200_OK_OR_ABORT(expr)                                                                         \
    do {                                                                                               \
        int rc_ = (expr);                                                                              \
        if (rc_ != 0) {                                                                                \
            b200_log("%s failed: %s (%s:%d)\n", #expr, b200_error_string(rc_), __FILE__, __LINE__);    \
            GGML_ABORT("ggml-b200: kernel library error");                                             \
        }                                                                                              \
    } while (0)

static bool is_hot_quant(enum ggml_type t) {
    return t == GGML_TYPE_Q4_0 || t == GGML_TYPE_Q8_0 || t == GGML_TYPE_Q4_K || t == GGML_TYPE_Q5_K || t == GGML_TYPE_Q6_K;
}
// .. and the formats the kernel library serves with its general kernels only (dequantize / GET_ROWS / CPY, one-warp-per-row mat-vec)
static bool is_block_quant(enum ggml_type t) {
    return is_hot_quant(t) || t == GGML_TYPE_Q4_1 || t == GGML_TYPE_Q5_0 || t == GGML_TYPE_Q5_1 || t == GGML_TYPE_IQ4_NL;
}

// ================================================================= contexts
struct b200_device_ctx {
    int         device;
    std::string name;         // "B200_0"
    std::string description;  // cudaDeviceProp::name
};

struct b200_buffer_ctx {
    int    device;
    void * base;
};

struct b200_backend_ctx;
static ggml_backend_buffer_type_t b200_buffer_type(int device);
static bool buft_is_b200(ggml_backend_buffer_type_t buft);
static bool buft_is_b200_split(ggml_backend_buffer_type_t buft);
// row-split weights (plugin_split.inc)
static bool split_mul_mat_supported(const struct ggml_tensor * op);
static void split_mul_mat(b200_backend_ctx * ctx, struct ggml_tensor * node);
static void split_prepare(b200_backend_ctx * ctx, const struct ggml_tensor * node);
static bool tensor_is_split(const struct ggml_tensor * t) { return t && t->buffer && buft_is_b200_split(t->buffer->buft); }

// ================================================================= device buffer
// mirrors the role of ggml_backend_cuda_buffer_* (src/ggml-cuda/ggml-cuda.cu:571-669)
static void b200_buffer_free(ggml_backend_buffer_t buffer) {
    b200_buffer_ctx * ctx = (b200_buffer_ctx *) buffer->context;
    cudaSetDevice(ctx->device);
    cudaFree(ctx->base);
    delete ctx;
}
static void * b200_buffer_get_base(ggml_backend_buffer_t buffer) {
    return ((b200_buffer_ctx *) buffer->context)->base;
}
static enum ggml_status b200_buffer_init_tensor(ggml_backend_buffer_t, struct ggml_tensor *) {
    return GGML_STATUS_SUCCESS;   // no padding contract: the kernels never read past a row's 16-byte ceiling
}
static void b200_buffer_memset_tensor(ggml_backend_buffer_t buffer, struct ggml_tensor * tensor, uint8_t value, size_t offset, size_t size) {
    b200_buffer_ctx * ctx = (b200_buffer_ctx *) buffer->context;
    CUDA_OK(cudaSetDevice(ctx->device));
    CUDA_OK(cudaMemsetAsync((char *) tensor->data + offset, value, size, cudaStreamPerThread));
    CUDA_OK(cudaStreamSynchronize(cudaStreamPerThread));
}
static void b200_buffer_set_tensor(ggml_backend_buffer_t buffer, struct ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    b200_buffer_ctx * ctx = (b200_buffer_ctx *) buffer->context;
    CUDA_OK(cudaSetDevice(ctx->device));
    CUDA_OK(cudaMemcpyAsync((char *) tensor->data + offset, data, size, cudaMemcpyHostToDevice, cudaStreamPerThread));
    CUDA_OK(cudaStreamSynchronize(cudaStreamPerThread));
}
static void b200_buffer_get_tensor(ggml_backend_buffer_t buffer, const struct ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    b200_buffer_ctx * ctx = (b200_buffer_ctx *) buffer->context;
    CUDA_OK(cudaSetDevice(ctx->device));
    CUDA_OK(cudaMemcpyAsync(data, (const char *) tensor->data + offset, size, cudaMemcpyDeviceToHost, cudaStreamPerThread));
    CUDA_OK(cudaStreamSynchronize(cudaStreamPerThread));
}
static bool b200_buffer_cpy_tensor(ggml_backend_buffer_t buffer, const struct ggml_tensor * src, struct ggml_tensor * dst) {
    if (!src->buffer || !buft_is_b200(src->buffer->buft)) return false;
    b200_buffer_ctx * sctx = (b200_buffer_ctx *) src->buffer->context;
    b200_buffer_ctx * dctx = (b200_buffer_ctx *) buffer->context;
    CUDA_OK(cudaSetDevice(dctx->device));
    if (sctx->device == dctx->device) {
        CUDA_OK(cudaMemcpyAsync(dst->data, src->data, ggml_nbytes(src), cudaMemcpyDeviceToDevice, cudaStreamPerThread));
    } else {
        CUDA_OK(cudaMemcpyPeerAsync(dst->data, dctx->device, src->data, sctx->device, ggml_nbytes(src), cudaStreamPerThread));
    }
    CUDA_OK(cudaStreamSynchronize(cudaStreamPerThread));
    return true;
}
static void b200_buffer_clear(ggml_backend_buffer_t buffer, uint8_t value) {
    b200_buffer_ctx * ctx = (b200_buffer_ctx *) buffer->context;
    CUDA_OK(cudaSetDevice(ctx->device));
    CUDA_OK(cudaMemset(ctx->base, value, buffer->size));
}
static const struct ggml_backend_buffer_i b200_buffer_iface = {
    /* .free_buffer   = */ b200_buffer_free,
    /* .get_base      = */ b200_buffer_get_base,
    /* .init_tensor   = */ b200_buffer_init_tensor,
    /* .memset_tensor = */ b200_buffer_memset_tensor,
    /* .set_tensor    = */ b200_buffer_set_tensor,
    /* .get_tensor    = */ b200_buffer_get_tensor,
    /* .cpy_tensor    = */ b200_buffer_cpy_tensor,
    /* .clear         = */ b200_buffer_clear,
    /* .reset         = */ NULL,
};

// ================================================================= device buffer type
// mirrors ggml_backend_cuda_buffer_type_* (src/ggml-cuda/ggml-cuda.cu:677-761)
struct b200_buft_ctx {
    int         device;
    std::string name;
};
static const char * b200_buft_get_name(ggml_backend_buffer_type_t buft) {
    return ((b200_buft_ctx *) buft->context)->name.c_str();
}
static ggml_backend_buffer_t b200_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    b200_buft_ctx * bctx = (b200_buft_ctx *) buft->context;
    if (cudaSetDevice(bctx->device) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    void * base = nullptr;
    // +256: bulk copies of a row may read up to the next 16-byte boundary past the last tensor
    cudaError_t e = cudaMalloc(&base, size + 256);
    if (e != cudaSuccess) {
        cudaGetLastError();   // recoverable, like the reference (ggml-cuda.cu:693-699)
        b200_log("allocating %.2f MiB on device %d failed: %s\n", size / 1048576.0, bctx->device, cudaGetErrorString(e));
        return nullptr;
    }
    b200_buffer_ctx * ctx = new b200_buffer_ctx{ bctx->device, base };
    return ggml_backend_buffer_init(buft, b200_buffer_iface, ctx, size);
}
static size_t b200_buft_get_alignment(ggml_backend_buffer_type_t) { return 128; }
static const struct ggml_backend_buffer_type_i b200_buft_iface = {
    /* .get_name       = */ b200_buft_get_name,
    /* .alloc_buffer   = */ b200_buft_alloc_buffer,
    /* .get_alignment  = */ b200_buft_get_alignment,
    /* .get_max_size   = */ NULL,
    /* .get_alloc_size = */ NULL,
    /* .is_host        = */ NULL,
};
static bool buft_is_b200(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == b200_buft_get_name; }

// ================================================================= pinned host buffer type
// same construction as ggml_backend_cuda_host_buffer_type (src/ggml-cuda/ggml-cuda.cu:1096-1160):
// a CPU buffer whose memory is page-locked, so set/get_tensor and the scheduler's copies run at
// full PCIe rate and can be asynchronous.
static const char * b200_host_buft_get_name(ggml_backend_buffer_type_t) { return "B200_Host"; }
static void b200_host_buffer_free(ggml_backend_buffer_t buffer) { cudaFreeHost(buffer->context); }
static ggml_backend_buffer_t b200_host_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    void * ptr = nullptr;
    if (cudaMallocHost(&ptr, size) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    ggml_backend_buffer_t buffer = ggml_backend_cpu_buffer_from_ptr(ptr, size);
    buffer->buft = buft;
    buffer->iface.free_buffer = b200_host_buffer_free;
    return buffer;
}

// ================================================================= backend (one stream)
struct graph_key_entry {
    int op; int type; const void * data;
    const void * src_data[3]; int src_type[3]; int src_usage[3];       // usage: weights buffers change the launch flags baked into a capture
    int32_t op_params[16];
    int64_t ne[4]; size_t nb[4];
    int64_t sne[3][4]; size_t snb[3][4];
    bool operator==(const graph_key_entry & o) const { return memcmp(this, &o, sizeof(*this)) == 0; }
};

struct b200_backend_ctx {
    int          device;
    std::string  name;
    cudaStream_t stream    = nullptr;
    void *       ws        = nullptr;
    size_t       ws_size   = 0;
    // CUDA-graph replay of repeated graph_compute() calls: a small LRU of captured graphs, so that the alternating
    // splits ggml_backend_sched hands a backend (one graph_compute per split) each find their own capture
    struct cached_graph {
        std::vector<graph_key_entry> key;
        cudaGraphExec_t exec = nullptr;      // captured on the second sight of the key, replayed from the third
        uint64_t last_use = 0;
    };
    static constexpr int GRAPH_CACHE = 8;
    cached_graph                 cache[GRAPH_CACHE];
    uint64_t                     use_clock = 0;
    bool                         graphs_enabled = true;
    uint64_t                     n_replays = 0, n_captures = 0, n_eager = 0;
    std::vector<b200_mul_mat_node> run;      // scratch: the current run of MUL_MAT nodes
};

static const char * b200_backend_get_name(ggml_backend_t backend) {
    return ((b200_backend_ctx *) backend->context)->name.c_str();
}
static void b200_backend_free(ggml_backend_t backend) {
    b200_backend_ctx * ctx = (b200_backend_ctx *) backend->context;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (auto & c : ctx->cache) if (c.exec) cudaGraphExecDestroy(c.exec);
    if (ctx->ws) cudaFree(ctx->ws);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    delete backend;
}
static void b200_backend_set_tensor_async(ggml_backend_t backend, struct ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    b200_backend_ctx * ctx = (b200_backend_ctx *) backend->context;
    CUDA_OK(cudaSetDevice(ctx->device));
    CUDA_OK(cudaMemcpyAsync((char *) tensor->data + offset, data, size, cudaMemcpyHostToDevice, ctx->stream));
}
static void b200_backend_get_tensor_async(ggml_backend_t backend, const struct ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    b200_backend_ctx * ctx = (b200_backend_ctx *) backend->context;
    CUDA_OK(cudaSetDevice(ctx->device));
    CUDA_OK(cudaMemcpyAsync(data, (const char *) tensor->data + offset, size, cudaMemcpyDeviceToHost, ctx->stream));
}
static void b200_backend_synchronize(ggml_backend_t backend) {
    b200_backend_ctx * ctx = (b200_backend_ctx *) backend->context;
    CUDA_OK(cudaSetDevice(ctx->device));
    CUDA_OK(cudaStreamSynchronize(ctx->stream));
    // bad DATA met by a kernel (a GET_ROWS index outside the table): the reference asserts on the host
    // (src/ggml-cpu/ops.cpp:5147); a kernel can only flag it, the abort happens here
    const unsigned flags = b200_device_error_flags(1);
    if (flags & 1u) {
        b200_log("GET_ROWS: row index outside the table (%s)\n", b200_error_string(B200_ERR_INDEX));
        GGML_ABORT("ggml-b200: GET_ROWS index out of range");
    }
}

static bool op_is_noop(const struct ggml_tensor * t) {
    return ggml_is_empty(t) || t->op == GGML_OP_NONE || t->op == GGML_OP_RESHAPE || t->op == GGML_OP_VIEW ||
           t->op == GGML_OP_PERMUTE || t->op == GGML_OP_TRANSPOSE;
}

// one node -> kernel-library calls on ctx->stream
static void b200_compute_node(b200_backend_ctx * ctx, struct ggml_tensor * node) {
    const struct ggml_tensor * s0 = node->src[0];
    const struct ggml_tensor * s1 = node->src[1];
    switch (node->op) {
        case GGML_OP_MUL_MAT:
            if (tensor_is_split(s0)) { split_mul_mat(ctx, node); break; }
            B200_OK_OR_ABORT(b200_mul_mat((int) s0->type, s0->data, s0->ne, s0->nb, (const float *) s1->data, s1->ne, s1->nb,
                                          (float *) node->data, ctx->ws, ctx->ws_size, ctx->stream));
            break;
        case GGML_OP_MUL_MAT_ID: {
            const struct ggml_tensor * ids = node->src[2];
            B200_OK_OR_ABORT(b200_mul_mat_id((int) s0->type, s0->data, s0->ne, s0->nb, (const float *) s1->data, s1->ne, s1->nb,
                                             (const int32_t *) ids->data, ids->ne, ids->nb, (float *) node->data, ctx->ws, ctx->ws_size, ctx->stream));
        } break;
        case GGML_OP_GET_ROWS: {
            // dst[:, i10, i11, i12] = dequant(src0[:, rows[i10, i11, i12], i11, i12])  (src/ggml-cpu/ops.cpp:5121-5165)
            const int out_type = node->type == GGML_TYPE_F32 ? B200_TYPE_F32 : B200_TYPE_F16;
            for (int64_t i12 = 0; i12 < s1->ne[2]; i12++) for (int64_t i11 = 0; i11 < s1->ne[1]; i11++) {
                const char * table = (const char *) s0->data + i11 * s0->nb[2] + i12 * s0->nb[3];
                const char * rows  = (const char *) s1->data + i11 * s1->nb[1] + i12 * s1->nb[2];
                char * dst = (char *) node->data + i11 * node->nb[2] + i12 * node->nb[3];
                B200_OK_OR_ABORT(b200_get_rows((int) s0->type, table, s0->nb[1], s0->ne[0], s0->ne[1], (const int32_t *) rows, s1->nb[0],
                                               s1->ne[0], dst, out_type, node->nb[1], ctx->stream));
            }
        } break;
        case GGML_OP_CPY:
        case GGML_OP_DUP:
        case GGML_OP_CONT: {
            const struct ggml_tensor * dst = node->op == GGML_OP_CPY ? node->src[1] : node;
            if (s0->type == GGML_TYPE_F32) {        // f32 -> Q8_0 / Q4_0 between tensors of one shape, dim 0 dense (supports_op)
                B200_OK_OR_ABORT(b200_quantize_rows_4d((int) dst->type, (const float *) s0->data, s0->ne, s0->nb, dst->data, dst->nb, ctx->stream));
                break;
            }
            const int out_type = dst->type == GGML_TYPE_F32 ? B200_TYPE_F32 : (dst->type == GGML_TYPE_F16 ? B200_TYPE_F16 : B200_TYPE_BF16);
            B200_OK_OR_ABORT(b200_dequantize_4d((int) s0->type, s0->data, s0->ne, s0->nb, dst->data, out_type, dst->ne, dst->nb, ctx->stream));
        } break;
        case GGML_OP_RMS_NORM: {
            float eps;
            memcpy(&eps, node->op_params, sizeof(float));
            B200_OK_OR_ABORT(b200_rms_norm((const float *) s0->data, s0->ne, s0->nb, (float *) node->data, eps, ctx->stream));
        } break;
        case GGML_OP_ADD:
        case GGML_OP_MUL:
            B200_OK_OR_ABORT(b200_bin_bcast(node->op == GGML_OP_ADD ? 0 : 1, (const float *) s0->data, s0->ne, s0->nb,
                                            (const float *) s1->data, s1->ne, s1->nb, (float *) node->data, ctx->stream));
            break;
        case GGML_OP_GLU: {
            // SWIGLU: one tensor whose rows hold [x | g] (or [g | x] when swapped), or two tensors (src/ggml-cpu/ops.cpp:3551-3610)
            const int32_t swapped = ((const int32_t *) node->op_params)[1];
            const int64_t nc = s1 ? s0->ne[0] : s0->ne[0] / 2;
            const char * x = (const char *) s0->data, * g = s1 ? (const char *) s1->data : (const char *) s0->data;
            if (!s1) { x += swapped ? nc * sizeof(float) : 0; g += swapped ? 0 : nc * sizeof(float); }
            B200_OK_OR_ABORT(b200_swiglu((const float *) x, s0->nb[1], (const float *) g, s1 ? s1->nb[1] : s0->nb[1], nc, ggml_nrows(s0),
                                         (float *) node->data, ctx->stream));
        } break;
        case GGML_OP_ROPE: {
            const int32_t * p = (const int32_t *) node->op_params;
            const int n_dims = p[1], mode = p[2], n_ctx_orig = p[4];
            float freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow, corr[2];
            memcpy(&freq_base, p + 5, 4); memcpy(&freq_scale, p + 6, 4); memcpy(&ext_factor, p + 7, 4);
            memcpy(&attn_factor, p + 8, 4); memcpy(&beta_fast, p + 9, 4); memcpy(&beta_slow, p + 10, 4);
            ggml_rope_yarn_corr_dims(n_dims, n_ctx_orig, freq_base, beta_fast, beta_slow, corr);
            const struct ggml_tensor * s2 = node->src[2];
            B200_OK_OR_ABORT(b200_rope((const float *) s0->data, s0->ne, s0->nb, (float *) node->data, node->nb, (const int32_t *) s1->data,
                                       s2 ? (const float *) s2->data : nullptr, n_dims, (mode & GGML_ROPE_TYPE_NEOX) ? 1 : 0,
                                       freq_base, freq_scale, ext_factor, attn_factor, corr[0], corr[1], ctx->stream));
        } break;
        case GGML_OP_SOFT_MAX: {
            float scale, max_bias;
            memcpy(&scale, (const float *) node->op_params + 0, 4); memcpy(&max_bias, (const float *) node->op_params + 1, 4);
            const struct ggml_tensor * s2 = node->src[2];
            const int64_t none[4] = { 0, 0, 1, 1 }; const size_t nonb[4] = { 0, 0, 0, 0 };
            B200_OK_OR_ABORT(b200_soft_max((const float *) s0->data, s0->ne, s0->nb, s1 ? s1->data : nullptr, s1 ? (int) s1->type : 0,
                                           s1 ? s1->ne : none, s1 ? s1->nb : nonb, s2 ? (const float *) s2->data : nullptr, scale, max_bias,
                                           (float *) node->data, ctx->stream));
        } break;
        default:
            b200_log("graph_compute: op %s was never claimed by supports_op\n", ggml_op_name(node->op));
            GGML_ABORT("ggml-b200: unsupported op");
    }
}

// A MUL_MAT node as the kernel library's list entry.  Weights living in a buffer that ggml marked
// GGML_BACKEND_BUFFER_USAGE_WEIGHTS are constant for the life of the graph.
static b200_mul_mat_node to_list_entry(const struct ggml_tensor * node) {
    const struct ggml_tensor * s0 = node->src[0];
    const struct ggml_tensor * s1 = node->src[1];
    b200_mul_mat_node e;
    memset(&e, 0, sizeof(e));
    e.type_a = (int32_t) s0->type;
    const struct ggml_tensor * base = s0->view_src ? s0->view_src : s0;
    e.flags = (base->buffer && base->buffer->usage == GGML_BACKEND_BUFFER_USAGE_WEIGHTS) ? B200_MUL_MAT_CONST_WEIGHTS : 0u;
    e.a_dev = s0->data; e.b_dev = (const float *) s1->data; e.dst_dev = (float *) node->data;
    for (int d = 0; d < 4; d++) { e.ne_a[d] = s0->ne[d]; e.nb_a[d] = s0->nb[d]; e.ne_b[d] = s1->ne[d]; e.nb_b[d] = s1->nb[d]; }
    return e;
}

// Runs of consecutive MUL_MAT nodes go to the library as ONE list (b200_mul_mat_batch): independent decode
// mat-vecs of a run (q/k/v, gate/up, or a whole per-layer GEMM graph) then share a persistent launch; the
// library cuts the list wherever a node depends on an earlier one.  fn_run(first, count) / fn_other(node).
template <typename FR, typename FO>
static void walk_graph(struct ggml_cgraph * cgraph, std::vector<b200_mul_mat_node> & run, FR fn_run, FO fn_other) {
    const int n = ggml_graph_n_nodes(cgraph);
    run.clear();
    for (int i = 0; i < n; i++) {
        struct ggml_tensor * node = ggml_graph_node(cgraph, i);
        if (op_is_noop(node)) continue;
        if (node->op == GGML_OP_MUL_MAT && !tensor_is_split(node->src[0])) { run.push_back(to_list_entry(node)); continue; }
        if (!run.empty()) { fn_run(run); run.clear(); }
        fn_other(node);
    }
    if (!run.empty()) { fn_run(run); run.clear(); }
}

static void make_key(struct ggml_cgraph * cgraph, std::vector<graph_key_entry> & key) {
    const int n = ggml_graph_n_nodes(cgraph);
    key.clear();
    key.reserve(n);
    for (int i = 0; i < n; i++) {
        const struct ggml_tensor * t = ggml_graph_node(cgraph, i);
        graph_key_entry e;
        memset(&e, 0, sizeof(e));
        e.op = (int) t->op; e.type = (int) t->type; e.data = t->data;
        memcpy(e.op_params, t->op_params, sizeof(e.op_params));
        for (int d = 0; d < 4; d++) { e.ne[d] = t->ne[d]; e.nb[d] = t->nb[d]; }
        for (int s = 0; s < 3; s++) {
            const struct ggml_tensor * src = t->src[s];
            if (!src) continue;
            e.src_data[s] = src->data; e.src_type[s] = (int) src->type;
            const struct ggml_tensor * base = src->view_src ? src->view_src : src;
            e.src_usage[s] = base->buffer ? (int) base->buffer->usage : 0;
            for (int d = 0; d < 4; d++) { e.sne[s][d] = src->ne[d]; e.snb[s][d] = src->nb[d]; }
        }
        key.push_back(e);
    }
}

static enum ggml_status b200_backend_graph_compute(ggml_backend_t backend, struct ggml_cgraph * cgraph) {
    b200_backend_ctx * ctx = (b200_backend_ctx *) backend->context;
    CUDA_OK(cudaSetDevice(ctx->device));

    // size the workspace before anything is launched (no allocation inside a capture)
    size_t need = 0;
    int n_work = 0;
    bool has_split = false;          // multi-device nodes are not captured into a CUDA graph
    std::vector<b200_mul_mat_node> & run = ctx->run;
    walk_graph(cgraph, run,
               [&](std::vector<b200_mul_mat_node> & r) {
                   n_work += (int) r.size();
                   const size_t w = b200_mul_mat_batch_workspace_size(r.data(), (int64_t) r.size());
                   if (w > need) need = w;
               },
               [&](struct ggml_tensor * node) {
                   n_work++;
                   if (tensor_is_split(node->src[0])) { has_split = true; split_prepare(ctx, node); }
                   if (node->op == GGML_OP_MUL_MAT_ID) {
                       const size_t w = b200_mul_mat_id_workspace_size((int) node->src[0]->type, node->src[0]->ne, node->src[1]->ne);
                       if (w > need) need = w;
                   }
               });
    if (need > ctx->ws_size) {
        CUDA_OK(cudaStreamSynchronize(ctx->stream));
        if (ctx->ws) CUDA_OK(cudaFree(ctx->ws));
        ctx->ws_size = need + (need >> 2) + (1u << 20);
        CUDA_OK(cudaMalloc(&ctx->ws, ctx->ws_size));
        for (auto & c : ctx->cache) {        // captured launches hold the old pointer
            if (c.exec) { cudaGraphExecDestroy(c.exec); c.exec = nullptr; }
            c.key.clear();
        }
    }

    // Replay policy: a graph seen twice in a row (same ops, shapes, strides, data pointers) is captured
    // on its second appearance and replayed from the third on.  Single-node graphs (test-backend-ops)
    // never pay for a capture.
    bool capture = false;
    b200_backend_ctx::cached_graph * slot = nullptr;
    if (ctx->graphs_enabled && n_work >= 4 && !has_split) {
        static thread_local std::vector<graph_key_entry> key;
        make_key(cgraph, key);
        b200_backend_ctx::cached_graph * lru = &ctx->cache[0];
        for (auto & c : ctx->cache) {
            if (!c.key.empty() && c.key == key) { slot = &c; break; }
            if (c.last_use < lru->last_use) lru = &c;
        }
        if (slot) {
            slot->last_use = ++ctx->use_clock;
            if (slot->exec) {
                CUDA_OK(cudaGraphLaunch(slot->exec, ctx->stream));
                ctx->n_replays++;
                return GGML_STATUS_SUCCESS;
            }
            capture = true;                  // second sight: capture now
        } else {                             // first sight: remember the key in the least recently used slot, run eagerly
            if (lru->exec) { cudaGraphExecDestroy(lru->exec); lru->exec = nullptr; }
            lru->key = key;
            lru->last_use = ++ctx->use_clock;
        }
    }
    if (capture) CUDA_OK(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed));
    walk_graph(cgraph, run,
               [&](std::vector<b200_mul_mat_node> & r) {
                   B200_OK_OR_ABORT(b200_mul_mat_batch(r.data(), (int64_t) r.size(), ctx->ws, ctx->ws_size, ctx->stream));
               },
               [&](struct ggml_tensor * node) { b200_compute_node(ctx, node); });
    if (capture) {
        cudaGraph_t graph = nullptr;
        CUDA_OK(cudaStreamEndCapture(ctx->stream, &graph));
        CUDA_OK(cudaGraphInstantiate(&slot->exec, graph, 0));
        CUDA_OK(cudaGraphDestroy(graph));
        ctx->n_captures++;
        CUDA_OK(cudaGraphLaunch(slot->exec, ctx->stream));
    } else {
        ctx->n_eager++;
    }
    return GGML_STATUS_SUCCESS;
}

static void b200_backend_event_record(ggml_backend_t backend, ggml_backend_event_t event) {
    b200_backend_ctx * ctx = (b200_backend_ctx *) backend->context;
    CUDA_OK(cudaEventRecord((cudaEvent_t) event->context, ctx->stream));
}
static void b200_backend_event_wait(ggml_backend_t backend, ggml_backend_event_t event) {
    b200_backend_ctx * ctx = (b200_backend_ctx *) backend->context;
    CUDA_OK(cudaStreamWaitEvent(ctx->stream, (cudaEvent_t) event->context, 0));
}

static bool b200_backend_cpy_tensor_async(ggml_backend_t backend_src, ggml_backend_t backend_dst, const struct ggml_tensor * src, struct ggml_tensor * dst);

static const struct ggml_backend_i b200_backend_iface = {
    /* .get_name           = */ b200_backend_get_name,
    /* .free               = */ b200_backend_free,
    /* .set_tensor_async   = */ b200_backend_set_tensor_async,
    /* .get_tensor_async   = */ b200_backend_get_tensor_async,
    /* .cpy_tensor_async   = */ b200_backend_cpy_tensor_async,
    /* .synchronize        = */ b200_backend_synchronize,
    /* .graph_plan_create  = */ NULL,
    /* .graph_plan_free    = */ NULL,
    /* .graph_plan_update  = */ NULL,
    /* .graph_plan_compute = */ NULL,
    /* .graph_compute      = */ b200_backend_graph_compute,
    /* .event_record       = */ b200_backend_event_record,
    /* .event_wait         = */ b200_backend_event_wait,
};

static ggml_guid_t b200_guid(void) {
    static ggml_guid guid = { 0xb2, 0x00, 0x51, 0x00, 0xa1, 0x7c, 0x4d, 0x05, 0x9e, 0x11, 0x67, 0x67, 0x6d, 0x6c, 0xb2, 0x00 };
    return &guid;
}

static bool b200_backend_cpy_tensor_async(ggml_backend_t backend_src, ggml_backend_t backend_dst, const struct ggml_tensor * src, struct ggml_tensor * dst) {
    // same-device copies between two b200 backends only; anything else goes through the generic path
    if (!ggml_guid_matches(backend_src->guid, b200_guid()) || !ggml_guid_matches(backend_dst->guid, b200_guid())) return false;
    if (!src->buffer || !dst->buffer || !buft_is_b200(src->buffer->buft) || !buft_is_b200(dst->buffer->buft)) return false;
    b200_backend_ctx * sctx = (b200_backend_ctx *) backend_src->context;
    b200_backend_ctx * dctx = (b200_backend_ctx *) backend_dst->context;
    if (sctx->device != dctx->device || backend_src != backend_dst) return false;
    CUDA_OK(cudaMemcpyAsync(dst->data, src->data, ggml_nbytes(dst), cudaMemcpyDeviceToDevice, sctx->stream));
    return true;
}

// ================================================================= device
// mirrors ggml_backend_cuda_device_* (src/ggml-cuda/ggml-cuda.cu:3153-3652)
static const char * b200_dev_get_name(ggml_backend_dev_t dev) { return ((b200_device_ctx *) dev->context)->name.c_str(); }
static const char * b200_dev_get_description(ggml_backend_dev_t dev) { return ((b200_device_ctx *) dev->context)->description.c_str(); }
static void b200_dev_get_memory(ggml_backend_dev_t dev, size_t * free, size_t * total) {
    b200_device_ctx * ctx = (b200_device_ctx *) dev->context;
    if (b200_device_info(ctx->device, nullptr, 0, free, total, nullptr, nullptr) != 0) { *free = 0; *total = 0; }
}
static enum ggml_backend_dev_type b200_dev_get_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }

static ggml_backend_buffer_type_t b200_host_buffer_type(void);

static void b200_dev_get_props(ggml_backend_dev_t dev, struct ggml_backend_dev_props * props) {
    props->name        = b200_dev_get_name(dev);
    props->description = b200_dev_get_description(dev);
    props->type        = GGML_BACKEND_DEVICE_TYPE_GPU;
    b200_dev_get_memory(dev, &props->memory_free, &props->memory_total);
    props->caps.async                = true;
    props->caps.host_buffer          = true;
    props->caps.buffer_from_host_ptr = false;
    props->caps.events               = true;
}

static ggml_backend_t b200_dev_init_backend(ggml_backend_dev_t dev, const char *) {
    b200_device_ctx * dctx = (b200_device_ctx *) dev->context;
    if (cudaSetDevice(dctx->device) != cudaSuccess) {
        cudaGetLastError();
        b200_log("cannot select device %d\n", dctx->device);
        return nullptr;
    }
    b200_backend_ctx * ctx = new b200_backend_ctx();
    ctx->device = dctx->device;
    ctx->name   = dctx->name;
    const char * g = getenv("GGML_B200_DISABLE_GRAPHS");
    ctx->graphs_enabled = !(g && *g && *g != '0');
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        cudaGetLastError();
        delete ctx;
        return nullptr;
    }
    ggml_backend_t backend = new ggml_backend{
        /* .guid    = */ b200_guid(),
        /* .iface   = */ b200_backend_iface,
        /* .device  = */ dev,
        /* .context = */ ctx,
    };
    return backend;
}

static ggml_backend_buffer_type_t b200_dev_get_buffer_type(ggml_backend_dev_t dev) {
    return b200_buffer_type(((b200_device_ctx *) dev->context)->device);
}
static ggml_backend_buffer_type_t b200_dev_get_host_buffer_type(ggml_backend_dev_t) { return b200_host_buffer_type(); }

static bool tensor_on_other_device(const struct ggml_tensor * t, int device) {
    if (!t || !t->buffer || !buft_is_b200(t->buffer->buft)) return false;
    return ((b200_buffer_ctx *) t->buffer->context)->device != device;
}

// The claim list.  Cheap, pure and thread-safe: it is called for every tensor (leafs included) by
// test-backend-ops (tests/test-backend-ops.cpp:1106-1115) and per node x backend by the scheduler.
static bool b200_dev_supports_op(ggml_backend_dev_t dev, const struct ggml_tensor * op) {
    const int device = ((b200_device_ctx *) dev->context)->device;
    for (int i = 0; i < GGML_MAX_SRC; i++) {
        if (tensor_on_other_device(op->src[i], device)) return false;
    }
    const struct ggml_tensor * s0 = op->src[0];
    const struct ggml_tensor * s1 = op->src[1];
    // row-split weights: only as src0 of a plain 2-D MUL_MAT (reference src/ggml-cuda/ggml-cuda.cu:3215-3222, 3276-3278)
    for (int i = 0; i < GGML_MAX_SRC; i++) {
        if (tensor_is_split(op->src[i])) return i == 0 && split_mul_mat_supported(op);
    }
    switch (op->op) {
        case GGML_OP_NONE:
        case GGML_OP_RESHAPE:
        case GGML_OP_VIEW:
        case GGML_OP_PERMUTE:
        case GGML_OP_TRANSPOSE:
            return true;
        case GGML_OP_MUL_MAT:
            return s0 && s1 && is_block_quant(s0->type) && s1->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 &&
                   s0->nb[0] == ggml_type_size(s0->type) && s1->nb[0] == sizeof(float) && ggml_is_contiguous(op) &&
                   s0->ne[2] > 0 && s0->ne[3] > 0 && s1->ne[2] % s0->ne[2] == 0 && s1->ne[3] % s0->ne[3] == 0 &&
                   s1->ne[2] <= 65535 && s1->ne[3] <= 65535;
        case GGML_OP_MUL_MAT_ID: {
            const struct ggml_tensor * ids = op->src[2];
            return s0 && s1 && ids && is_block_quant(s0->type) && s1->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && ids->type == GGML_TYPE_I32 &&
                   s0->nb[0] == ggml_type_size(s0->type) && s1->nb[0] == sizeof(float) && ggml_is_contiguous(op) && s0->ne[3] == 1 && s1->ne[3] == 1 &&
                   s1->ne[1] >= 1 && ids->ne[0] % s1->ne[1] == 0 && ids->ne[0] <= 65535 && ids->ne[1] <= 65535;
        }
        case GGML_OP_GET_ROWS:
            return s0 && s1 && is_block_quant(s0->type) && s1->type == GGML_TYPE_I32 && op->type == GGML_TYPE_F32 &&
                   s0->nb[0] == ggml_type_size(s0->type) && op->nb[0] == sizeof(float) && s1->ne[3] == 1;
        case GGML_OP_CPY:
        case GGML_OP_DUP:
        case GGML_OP_CONT: {
            const struct ggml_tensor * dst = op->op == GGML_OP_CPY ? s1 : op;
            if (s0 && dst && s0->type == GGML_TYPE_F32 && (dst->type == GGML_TYPE_Q8_0 || dst->type == GGML_TYPE_Q4_0))
                return op->op == GGML_OP_CPY && ggml_are_same_shape(s0, dst) && s0->nb[0] == sizeof(float) && dst->nb[0] == ggml_type_size(dst->type) &&
                       s0->ne[0] % 32 == 0;
            return s0 && dst && is_block_quant(s0->type) && (dst->type == GGML_TYPE_F32 || dst->type == GGML_TYPE_F16 || dst->type == GGML_TYPE_BF16) &&
                   s0->nb[0] == ggml_type_size(s0->type) && ggml_nelements(s0) == ggml_nelements(dst);
        }
        // the f32 ops between the MUL_MATs of a decoder layer (csrc/layer_ops.cu)
        case GGML_OP_RMS_NORM:
            return s0 && s0->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && s0->nb[0] == sizeof(float) && ggml_is_contiguous(op);
        case GGML_OP_ADD:
        case GGML_OP_MUL:
            return s0 && s1 && s0->type == GGML_TYPE_F32 && s1->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 &&
                   ggml_is_contiguous(op) && ggml_are_same_shape(s0, op) && ggml_can_repeat(s1, s0);
        case GGML_OP_GLU:
            return s0 && ggml_get_glu_op(op) == GGML_GLU_OP_SWIGLU && s0->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 &&
                   ggml_is_contiguous_1(s0) && ggml_is_contiguous(op) && (!s1 || (s1->type == GGML_TYPE_F32 && ggml_is_contiguous_1(s1)));
        case GGML_OP_ROPE: {
            const int mode = ((const int32_t *) op->op_params)[2];
            const struct ggml_tensor * s2 = op->src[2];
            return s0 && s1 && s0->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && s1->type == GGML_TYPE_I32 &&
                   (mode & ~GGML_ROPE_TYPE_NEOX) == 0 && s0->nb[0] == sizeof(float) && op->nb[0] == sizeof(float) && s0->ne[0] % 2 == 0 &&
                   (!s2 || s2->type == GGML_TYPE_F32);
        }
        case GGML_OP_SOFT_MAX: {
            const struct ggml_tensor * s2 = op->src[2];
            return s0 && s0->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && s0->nb[0] == sizeof(float) && ggml_is_contiguous(op) &&
                   (!s1 || ((s1->type == GGML_TYPE_F32 || s1->type == GGML_TYPE_F16) && s1->nb[0] == ggml_type_size(s1->type))) &&
                   (!s2 || s2->type == GGML_TYPE_F32);
        }
        default:
            return false;
    }
}

static bool b200_dev_supports_buft(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft) {
    if (buft_is_b200_split(buft)) return true;       // any device of the box can be the main device of a split MUL_MAT
    if (!buft_is_b200(buft)) return false;
    return ((b200_buft_ctx *) buft->context)->device == ((b200_device_ctx *) dev->context)->device;
}

// offload weights that live in host memory when the batch is large enough to amortise the copy
// (the reference's rule: batch >= 32, src/ggml-cuda/ggml-cuda.cu:3583-3604)
static bool b200_dev_offload_op(ggml_backend_dev_t, const struct ggml_tensor * op) {
    return op->op == GGML_OP_MUL_MAT && op->ne[1] >= 32;
}

static ggml_backend_event_t b200_dev_event_new(ggml_backend_dev_t dev) {
    b200_device_ctx * ctx = (b200_device_ctx *) dev->context;
    CUDA_OK(cudaSetDevice(ctx->device));
    cudaEvent_t ev;
    CUDA_OK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    return new ggml_backend_event{ /* .device = */ dev, /* .context = */ ev };
}
static void b200_dev_event_free(ggml_backend_dev_t, ggml_backend_event_t event) {
    cudaEventDestroy((cudaEvent_t) event->context);
    delete event;
}
static void b200_dev_event_synchronize(ggml_backend_dev_t, ggml_backend_event_t event) {
    CUDA_OK(cudaEventSynchronize((cudaEvent_t) event->context));
}

static const struct ggml_backend_device_i b200_device_iface = {
    /* .get_name             = */ b200_dev_get_name,
    /* .get_description      = */ b200_dev_get_description,
    /* .get_memory           = */ b200_dev_get_memory,
    /* .get_type             = */ b200_dev_get_type,
    /* .get_props            = */ b200_dev_get_props,
    /* .init_backend         = */ b200_dev_init_backend,
    /* .get_buffer_type      = */ b200_dev_get_buffer_type,
    /* .get_host_buffer_type = */ b200_dev_get_host_buffer_type,
    /* .buffer_from_host_ptr = */ NULL,
    /* .supports_op          = */ b200_dev_supports_op,
    /* .supports_buft        = */ b200_dev_supports_buft,
    /* .offload_op           = */ b200_dev_offload_op,
    /* .event_new            = */ b200_dev_event_new,
    /* .event_free           = */ b200_dev_event_free,
    /* .event_synchronize    = */ b200_dev_event_synchronize,
};

// ================================================================= registry + singletons
struct b200_reg_ctx {
    std::vector<ggml_backend_dev_t> devices;
};

static const char * b200_reg_get_name(ggml_backend_reg_t) { return "B200"; }
static size_t b200_reg_get_device_count(ggml_backend_reg_t reg) { return ((b200_reg_ctx *) reg->context)->devices.size(); }
static ggml_backend_dev_t b200_reg_get_device(ggml_backend_reg_t reg, size_t index) {
    b200_reg_ctx * ctx = (b200_reg_ctx *) reg->context;
    GGML_ASSERT(index < ctx->devices.size());
    return ctx->devices[index];
}

static struct ggml_backend_feature * b200_get_features(ggml_backend_reg_t) {
    static struct ggml_backend_feature features[] = {
        { "ARCH", "sm_100a" }, { "TMA_MMVQ", "1" }, { "TCGEN05_MMQ", "1" }, { "CUDA_GRAPHS", "1" }, { nullptr, nullptr },
    };
    return features;
}

static void * b200_reg_get_proc_address(ggml_backend_reg_t, const char * name) {
    if (strcmp(name, "ggml_backend_split_buffer_type") == 0) return (void *) ggml_backend_b200_split_buffer_type;
    if (strcmp(name, "ggml_backend_get_features") == 0) return (void *) b200_get_features;
    if (strcmp(name, "ggml_backend_b200_load_gguf") == 0) return (void *) ggml_backend_b200_load_gguf;
    return nullptr;
}

static const struct ggml_backend_reg_i b200_reg_iface = {
    /* .get_name         = */ b200_reg_get_name,
    /* .get_device_count = */ b200_reg_get_device_count,
    /* .get_device       = */ b200_reg_get_device,
    /* .get_proc_address = */ b200_reg_get_proc_address,
};

ggml_backend_reg_t ggml_backend_b200_reg(void) {
    static ggml_backend_reg reg;
    static bool initialized = false;
    static std::mutex mutex;
    std::lock_guard<std::mutex> lock(mutex);
    if (!initialized) {
        b200_reg_ctx * ctx = new b200_reg_ctx;
        const int n = b200_device_count();
        for (int i = 0; i < n && i < B200_MAX_DEVICES; i++) {
            char name[256] = { 0 };
            b200_device_info(i, name, sizeof(name), nullptr, nullptr, nullptr, nullptr);
            b200_device_ctx * dctx = new b200_device_ctx{ i, "B200_" + std::to_string(i), name };
            ctx->devices.push_back(new ggml_backend_device{ /* .iface = */ b200_device_iface, /* .reg = */ &reg, /* .context = */ dctx });
        }
        reg = ggml_backend_reg{ /* .api_version = */ GGML_BACKEND_API_VERSION, /* .iface = */ b200_reg_iface, /* .context = */ ctx };
        initialized = true;
    }
    return &reg;
}

static ggml_backend_buffer_type_t b200_buffer_type(int device) {
    static struct ggml_backend_buffer_type bufts[B200_MAX_DEVICES];
    static bool initialized = false;
    static std::mutex mutex;
    std::lock_guard<std::mutex> lock(mutex);
    if (device < 0 || device >= B200_MAX_DEVICES) return nullptr;
    if (!initialized) {
        ggml_backend_reg_t reg = ggml_backend_b200_reg();
        const int n = (int) b200_reg_get_device_count(reg);
        for (int i = 0; i < B200_MAX_DEVICES; i++) {
            bufts[i] = ggml_backend_buffer_type{
                /* .iface   = */ b200_buft_iface,
                /* .device  = */ i < n ? b200_reg_get_device(reg, i) : nullptr,
                /* .context = */ new b200_buft_ctx{ i, "B200_" + std::to_string(i) },
            };
        }
        initialized = true;
    }
    return &bufts[device];
}

static ggml_backend_buffer_type_t b200_host_buffer_type(void) {
    static struct ggml_backend_buffer_type buft;
    static bool initialized = false;
    static std::mutex mutex;
    std::lock_guard<std::mutex> lock(mutex);
    if (!initialized) {
        ggml_backend_reg_t reg = ggml_backend_b200_reg();
        buft.iface = ggml_backend_cpu_buffer_type()->iface;     // is_host = true, alignment, alloc size as CPU
        buft.iface.get_name     = b200_host_buft_get_name;
        buft.iface.alloc_buffer = b200_host_buft_alloc_buffer;
        buft.device  = b200_reg_get_device_count(reg) > 0 ? b200_reg_get_device(reg, 0) : nullptr;
        buft.context = nullptr;
        initialized = true;
    }
    return &buft;
}

// ================================================================= exported C surface
extern "C" {

ggml_backend_reg_t ggml_backend_init(void) { return ggml_backend_b200_reg(); }
int ggml_backend_score(void) { return b200_device_count() > 0 ? 100 : 0; }

ggml_backend_t ggml_backend_b200_init(int device) {
    ggml_backend_reg_t reg = ggml_backend_b200_reg();
    if (device < 0 || (size_t) device >= b200_reg_get_device_count(reg)) {
        b200_log("invalid device %d\n", device);
        return nullptr;
    }
    return b200_dev_init_backend(b200_reg_get_device(reg, device), nullptr);
}
int ggml_backend_is_b200(ggml_backend_t backend) { return backend != nullptr && ggml_guid_matches(backend->guid, b200_guid()); }
ggml_backend_buffer_type_t ggml_backend_b200_buffer_type(int device) { return b200_buffer_type(device); }
ggml_backend_buffer_type_t ggml_backend_b200_host_buffer_type(void) { return b200_host_buffer_type(); }
int  ggml_backend_b200_get_device_count(void) { return b200_device_count(); }
void ggml_backend_b200_get_device_description(int device, char * description, size_t description_size) {
    if (b200_device_info(device, description, description_size, nullptr, nullptr, nullptr, nullptr) != 0 && description_size) description[0] = 0;
}
void ggml_backend_b200_get_device_memory(int device, size_t * free_bytes, size_t * total_bytes) {
    if (b200_device_info(device, nullptr, 0, free_bytes, total_bytes, nullptr, nullptr) != 0) { *free_bytes = 0; *total_bytes = 0; }
}

// GGUF file -> tensors in a device buffer, in one call: what gguf_init_from_file(params.ctx != NULL) does into host memory
// (src/gguf.cpp:640-735: one ggml_context holding every tensor of the file, named, data filled in), here with the data in
// a buffer of ggml_backend_b200_buffer_type(device), streamed through the kernel library's pinned staging ring
// (b200_gguf_load_tensors) instead of read() + ggml_backend_tensor_set per tensor.  *ctx_out owns the tensor metadata
// (ggml_free), the returned buffer owns the bytes (ggml_backend_buffer_free); NULL on any failure.
ggml_backend_buffer_t ggml_backend_b200_load_gguf(const char * fname, int device, struct ggml_context ** ctx_out) {
    if (!fname || !ctx_out) return nullptr;
    *ctx_out = nullptr;
    ggml_backend_buffer_type_t buft = b200_buffer_type(device);
    if (!buft || !buft->device) { b200_log("invalid device %d\n", device); return nullptr; }
    b200_gguf * g = nullptr;
    if (b200_gguf_open(fname, &g) != 0) { b200_log("%s: %s\n", fname, b200_gguf_last_error()); return nullptr; }
    const int64_t n = b200_gguf_n_tensors(g);
    struct ggml_init_params ip = { /* .mem_size = */ (size_t) (n + 1) * ggml_tensor_overhead(), /* .mem_buffer = */ nullptr, /* .no_alloc = */ true };
    struct ggml_context * ctx = ggml_init(ip);
    std::vector<int64_t> ids;
    std::vector<void *>  dsts;
    ggml_backend_buffer_t buf = nullptr;
    bool ok = ctx != nullptr;
    for (int64_t i = 0; ok && i < n; i++) {
        b200_gguf_tensor_info ti;
        ok = b200_gguf_tensor(g, i, &ti) == 0;
        if (!ok) break;
        struct ggml_tensor * t = ggml_new_tensor(ctx, (enum ggml_type) ti.type, GGML_MAX_DIMS, ti.ne);
        ok = t != nullptr && ggml_nbytes(t) == ti.nbytes;
        if (ok) ggml_set_name(t, ti.name);
    }
    if (ok && n > 0) {
        buf = ggml_backend_alloc_ctx_tensors_from_buft(ctx, buft);
        ok = buf != nullptr;
    }
    if (ok && buf) {
        ggml_backend_buffer_set_usage(buf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
        int64_t i = 0;
        for (struct ggml_tensor * t = ggml_get_first_tensor(ctx); t != nullptr; t = ggml_get_next_tensor(ctx, t), i++) {
            ids.push_back(i);
            dsts.push_back(t->data);
        }
        int prev = 0;
        CUDA_OK(cudaGetDevice(&prev));
        CUDA_OK(cudaSetDevice(device));
        const int rc = b200_gguf_load_tensors(g, ids.data(), dsts.data(), (int64_t) ids.size(), /* stream = */ nullptr);
        if (rc == 0) CUDA_OK(cudaStreamSynchronize(nullptr));
        CUDA_OK(cudaSetDevice(prev));
        if (rc != 0) { b200_log("%s: %s\n", fname, rc == B200_ERR_FILE ? b200_gguf_last_error() : b200_error_string(rc)); ok = false; }
    }
    b200_gguf_close(g);
    if (!ok) {
        if (buf) ggml_backend_buffer_free(buf);
        if (ctx) ggml_free(ctx);
        return nullptr;
    }
    *ctx_out = ctx;
    return buf;
}

} // extern "C"

#include "plugin_split.inc"
