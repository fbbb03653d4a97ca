/*
 * ggml_b200_plugin.h — what libggml-b200.so exports.  The plugin is an out-of-tree ggml backend:
 * an UNMODIFIED ggml loads it with ggml_backend_load(path) / GGML_BACKEND_PATH
 * (reference src/ggml-backend-reg.cpp:239-283, 595-599), which dlsym()s the two loader symbols below
 * (typedefs: reference src/ggml-backend-impl.h:214-218).  The vtables it fills are the reference's
 * own structs from src/ggml-backend-impl.h (the plugin is compiled against those headers); they
 * replace the ones in src/ggml-cuda/ggml-cuda.cu:571-761 (buffer + buffer type), :2530-3082
 * (backend), :3153-3652 (device), :3660-3753 (registry).
 *
 * The named-function surface mirrors include/ggml-cuda.h:22-44 with the b200 prefix so an example
 * that says ggml_backend_cuda_init(0) can say ggml_backend_b200_init(0) instead.
 */
#ifndef GGML_B200_PLUGIN_H
#define GGML_B200_PLUGIN_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef B200_API
#define B200_API __attribute__((visibility("default")))
#endif

/* opaque ggml handles (include/ggml-backend.h:19-24) */
struct ggml_backend;
struct ggml_backend_reg;
struct ggml_backend_buffer_type;
struct ggml_backend_buffer;
struct ggml_context;

/* loader entry points (src/ggml-backend-impl.h:214-218) */
B200_API struct ggml_backend_reg * ggml_backend_init(void);
B200_API int                       ggml_backend_score(void);   /* 0 when no sm_100 device is present */

/* include/ggml-cuda.h:22-44 analogues */
B200_API struct ggml_backend_reg *         ggml_backend_b200_reg(void);
B200_API struct ggml_backend *             ggml_backend_b200_init(int device);
B200_API int                               ggml_backend_is_b200(struct ggml_backend * backend);
B200_API struct ggml_backend_buffer_type * ggml_backend_b200_buffer_type(int device);
B200_API struct ggml_backend_buffer_type * ggml_backend_b200_split_buffer_type(int main_device, const float * tensor_split);
B200_API struct ggml_backend_buffer_type * ggml_backend_b200_host_buffer_type(void);
B200_API int                               ggml_backend_b200_get_device_count(void);
B200_API void                              ggml_backend_b200_get_device_description(int device, char * description, size_t description_size);
B200_API void                              ggml_backend_b200_get_device_memory(int device, size_t * free_bytes, size_t * total_bytes);


/* GGUF file -> named tensors in a buffer of ggml_backend_b200_buffer_type(device), in one call (also reachable as the
 * registry proc "ggml_backend_b200_load_gguf").  The device-side counterpart of gguf_init_from_file with params.ctx set
 * (include/gguf.h:72-80, src/gguf.cpp:640-735), which fills a host context: *ctx_out receives a no_alloc ggml_context
 * holding every tensor of the file (ggml_get_tensor(ctx, name)), their data streamed file -> pinned staging ring -> HBM
 * by b200_gguf_load_tensors (include/ggml_b200.h) and marked GGML_BACKEND_BUFFER_USAGE_WEIGHTS.  Key/value metadata:
 * b200_gguf_* or the reference's own gguf_init_from_file(no_alloc).  Returns NULL (and logs why) on failure; a file
 * without tensors returns NULL with *ctx_out set. */
B200_API struct ggml_backend_buffer * ggml_backend_b200_load_gguf(const char * fname, int device, struct ggml_context ** ctx_out);

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_PLUGIN_H */
