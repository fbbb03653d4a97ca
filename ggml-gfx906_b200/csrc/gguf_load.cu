// gguf_load.cu — GGUF file -> HBM: the data format on the input side of the hot path.
//
// Two halves, both behind the C ABI of include/ggml_b200.h:
//   * a reader of the GGUF container with the acceptance rules of the reference's gguf_init_from_file_impl
//     (src/gguf.cpp:319-735; format: include/gguf.h:1-31): magic, version 2..3 (an endian-swapped version is named as
//     such), key/value pairs of the 13 value types with duplicate keys refused, general.alignment a power of two, tensor
//     infos (unique names shorter than GGML_MAX_NAME, <= 4 non-negative extents whose product fits int64, a defined
//     type whose block size divides ne[0]) and data offsets that are exactly the running, alignment-padded sum of the
//     tensor sizes.  Host only: it runs (and is tested) without a GPU.
//   * the loader: tensor bytes go file -> pinned staging ring -> device with cudaMemcpyAsync, one reader thread per
//     stage (pread) filling the next stages while earlier ones are in flight over PCIe; nothing is ever staged in pageable memory and no
//     per-tensor allocation happens.  Device image: tensor i at b200_gguf_device_offset(i), every tensor aligned to
//     256 bytes (TMA boxes and 16-byte bulk copies of the kernels need 16; the buffer type advertises 128).  The bytes
//     are ggml's canonical block bytes, unchanged: what ggml_backend_tensor_get must hand back.
//     replaces the read + ggml_backend_tensor_set loop the reference's examples run per tensor
//     (ggml_backend_cuda_buffer_set_tensor: a synchronous pageable cudaMemcpy each, src/ggml-cuda/ggml-cuda.cu:614-621).
#include "common.cuh"

#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <fcntl.h>
#include <mutex>
#include <string>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>
#include <vector>

namespace {

// ---------------------------------------------------------------- ggml type table (every type a GGUF file may name)
// (elements per block, bytes per block) for enum ggml_type 0..39 (include/ggml.h:376-417, sizes from the block structs of
// src/ggml-common.h); {0, 0} = removed / undefined ids, which the reference rejects through blck_size == 0 (gguf.cpp:588).
struct tinfo { int qk, bytes; const char * name; };
const tinfo g_types[40] = {
    /* 0 f32     */ {1, 4, "f32"},      /* 1 f16     */ {1, 2, "f16"},      /* 2 q4_0    */ {32, 18, "q4_0"},   /* 3 q4_1    */ {32, 20, "q4_1"},
    /* 4         */ {0, 0, "removed"},  /* 5         */ {0, 0, "removed"},  /* 6 q5_0    */ {32, 22, "q5_0"},   /* 7 q5_1    */ {32, 24, "q5_1"},
    /* 8 q8_0    */ {32, 34, "q8_0"},   /* 9 q8_1    */ {32, 36, "q8_1"},   /* 10 q2_K   */ {256, 84, "q2_K"},  /* 11 q3_K   */ {256, 110, "q3_K"},
    /* 12 q4_K   */ {256, 144, "q4_K"}, /* 13 q5_K   */ {256, 176, "q5_K"}, /* 14 q6_K   */ {256, 210, "q6_K"}, /* 15 q8_K   */ {256, 292, "q8_K"},
    /* 16 iq2_xxs*/ {256, 66, "iq2_xxs"}, /* 17 iq2_xs */ {256, 74, "iq2_xs"}, /* 18 iq3_xxs*/ {256, 98, "iq3_xxs"}, /* 19 iq1_s  */ {256, 50, "iq1_s"},
    /* 20 iq4_nl */ {32, 18, "iq4_nl"}, /* 21 iq3_s  */ {256, 110, "iq3_s"}, /* 22 iq2_s  */ {256, 82, "iq2_s"}, /* 23 iq4_xs */ {256, 136, "iq4_xs"},
    /* 24 i8     */ {1, 1, "i8"},       /* 25 i16    */ {1, 2, "i16"},      /* 26 i32    */ {1, 4, "i32"},      /* 27 i64    */ {1, 8, "i64"},
    /* 28 f64    */ {1, 8, "f64"},      /* 29 iq1_m  */ {256, 56, "iq1_m"}, /* 30 bf16   */ {1, 2, "bf16"},     /* 31        */ {0, 0, "removed"},
    /* 32        */ {0, 0, "removed"},  /* 33        */ {0, 0, "removed"},  /* 34 tq1_0  */ {256, 54, "tq1_0"}, /* 35 tq2_0  */ {256, 66, "tq2_0"},
    /* 36        */ {0, 0, "removed"},  /* 37        */ {0, 0, "removed"},  /* 38        */ {0, 0, "removed"},  /* 39 mxfp4  */ {32, 17, "mxfp4"},
};
constexpr int    MAX_NAME     = 64;          // GGML_MAX_NAME (include/ggml.h:228)
constexpr size_t DEV_ALIGN    = 256;
constexpr size_t STAGE_BYTES  = 16u << 20;   // one staging buffer; 8 of them = 128 MiB of pinned host memory per open loader
constexpr int    N_STAGES     = 8;           // = reader threads: one pread stream per stage (a single one tops out near 8 GB/s)

enum kv_type : int32_t { KV_U8, KV_I8, KV_U16, KV_I16, KV_U32, KV_I32, KV_F32, KV_BOOL, KV_STR, KV_ARR, KV_U64, KV_I64, KV_F64, KV_COUNT };
const int kv_size[KV_COUNT] = {1, 1, 2, 2, 4, 4, 4, 1, 0, 0, 8, 8, 8};

struct kv_entry {
    std::string              key;
    int32_t                  type = -1;      // element type (never KV_ARR)
    bool                     is_array = false;
    uint64_t                 n = 1;
    std::vector<uint8_t>     data;           // scalars / arrays of fixed-size elements, file byte order (little endian)
    std::vector<std::string> strs;           // string value / array of strings
};

struct tensor_entry {
    std::string name;
    int32_t     type = 0, n_dims = 0;
    int64_t     ne[4] = {1, 1, 1, 1};
    uint64_t    offset = 0;                  // relative to the data section
    uint64_t    nbytes = 0;
    uint64_t    dev_offset = 0;
};

thread_local std::string t_error;
void set_error(const char * fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    t_error = buf;
}

struct file_reader {
    FILE * f;
    template <typename T> bool rd(T & v) const { return fread(&v, 1, sizeof(T), f) == sizeof(T); }
    bool rd_str(std::string & s, uint64_t limit) const {
        uint64_t n = 0;
        if (!rd(n) || n > limit) return false;   // a length beyond the file is a truncated / corrupt file, not an allocation request
        s.resize((size_t) n);
        return n == 0 || fread(&s[0], 1, (size_t) n, f) == n;
    }
};

struct piece { size_t stage_off; char * dst; size_t n; };

} // namespace

struct b200_gguf {
    std::string               path;
    int                       fd = -1;
    uint64_t                  file_bytes = 0;
    uint32_t                  version = 0;
    size_t                    alignment = 32;
    uint64_t                  data_offset = 0, data_bytes = 0, device_bytes = 0;
    std::vector<kv_entry>     kv;
    std::vector<tensor_entry> tensors;
    // staging ring (allocated by the first load on a device, released by close)
    char *                    stage[N_STAGES] = {};
    cudaEvent_t               ev[N_STAGES] = {};
    bool                      ev_pending[N_STAGES] = {};
};

namespace {

bool parse(b200_gguf & g, FILE * f) {
    file_reader r{f};
    char magic[4];
    if (fread(magic, 1, 4, f) != 4) { set_error("failed to read magic"); return false; }
    if (memcmp(magic, "GGUF", 4) != 0) { set_error("invalid magic characters, expected 'GGUF'"); return false; }
    if (!r.rd(g.version)) { set_error("failed to read header"); return false; }
    if (g.version == 0) { set_error("bad GGUF version: 0"); return false; }
    if ((g.version & 0xFFFFu) == 0) { set_error("GGUF version %u: mismatch between the host and file endianness?", g.version); return false; }
    if (g.version == 1) { set_error("GGUFv1 is no longer supported"); return false; }
    if (g.version > 3) { set_error("GGUF version %u, this reader supports up to 3", g.version); return false; }
    int64_t n_tensors = 0, n_kv = 0;
    if (!r.rd(n_tensors) || !r.rd(n_kv)) { set_error("failed to read header"); return false; }
    // every entry takes at least 8 bytes of file: a count beyond that is corrupt (the reference bounds it by SIZE_MAX / sizeof)
    if (n_tensors < 0 || (uint64_t) n_tensors > g.file_bytes / 8) { set_error("number of tensors is %lld", (long long) n_tensors); return false; }
    if (n_kv < 0 || (uint64_t) n_kv > g.file_bytes / 8) { set_error("number of key value pairs is %lld", (long long) n_kv); return false; }

    for (int64_t i = 0; i < n_kv; i++) {
        kv_entry e;
        if (!r.rd_str(e.key, g.file_bytes)) { set_error("failed to read key %lld", (long long) i); return false; }
        for (const kv_entry & o : g.kv) if (o.key == e.key) { set_error("duplicate key '%s'", e.key.c_str()); return false; }
        if (!r.rd(e.type)) { set_error("failed to read key-value pairs"); return false; }
        if (e.type == KV_ARR) {
            e.is_array = true;
            if (!r.rd(e.type) || !r.rd(e.n)) { set_error("failed to read key-value pairs"); return false; }
        }
        if (e.type < 0 || e.type >= KV_COUNT || e.type == KV_ARR) { set_error("key '%s' has invalid GGUF type %d", e.key.c_str(), e.type); return false; }
        if (e.type == KV_STR) {
            if (e.n > g.file_bytes / 8) { set_error("key '%s': array length beyond the file", e.key.c_str()); return false; }
            e.strs.resize((size_t) e.n);
            for (std::string & s : e.strs) if (!r.rd_str(s, g.file_bytes)) { set_error("failed to read value for key '%s'", e.key.c_str()); return false; }
        } else {
            const uint64_t sz = (uint64_t) kv_size[e.type];
            if (e.n > g.file_bytes / sz) { set_error("key '%s': array length beyond the file", e.key.c_str()); return false; }
            e.data.resize((size_t) (e.n * sz));
            if (!e.data.empty() && fread(e.data.data(), 1, e.data.size(), f) != e.data.size()) { set_error("failed to read value for key '%s'", e.key.c_str()); return false; }
        }
        g.kv.push_back(std::move(e));
    }
    g.alignment = 32;   // GGUF_DEFAULT_ALIGNMENT
    for (const kv_entry & e : g.kv) if (e.key == "general.alignment") {
        if (e.is_array || e.type != KV_U32) { set_error("general.alignment is not a uint32"); return false; }   // the reference aborts on the wrong type
        uint32_t a; memcpy(&a, e.data.data(), 4); g.alignment = a;
    }
    if (g.alignment == 0 || (g.alignment & (g.alignment - 1)) != 0) { set_error("alignment %zu is not a power of 2", g.alignment); return false; }

    for (int64_t i = 0; i < n_tensors; i++) {
        tensor_entry t;
        if (!r.rd_str(t.name, g.file_bytes)) { set_error("failed to read tensor name %lld", (long long) i); return false; }
        if (t.name.size() >= (size_t) MAX_NAME) { set_error("tensor name %lld is too long: %zu >= %d", (long long) i, t.name.size(), MAX_NAME); return false; }
        for (const tensor_entry & o : g.tensors) if (o.name == t.name) { set_error("duplicate tensor name '%s'", t.name.c_str()); return false; }
        uint32_t n_dims = 0;
        if (!r.rd(n_dims)) { set_error("failed to read tensor info"); return false; }
        if (n_dims > 4) { set_error("tensor '%s' has invalid number of dimensions: %u > 4", t.name.c_str(), n_dims); return false; }
        t.n_dims = (int32_t) n_dims;
        for (uint32_t j = 0; j < n_dims; j++) {
            if (!r.rd(t.ne[j])) { set_error("failed to read tensor info"); return false; }
            if (t.ne[j] < 0) { set_error("tensor '%s' dimension %u has invalid number of elements: %lld < 0", t.name.c_str(), j, (long long) t.ne[j]); return false; }
        }
        // the element count must be representable (same three comparisons as gguf.cpp:561-563; a zero extent divides by zero
        // there, so it is refused here)
        if (t.ne[1] == 0 || t.ne[2] == 0 || t.ne[3] == 0 ||
            INT64_MAX / t.ne[1] <= t.ne[0] || INT64_MAX / t.ne[2] <= t.ne[0] * t.ne[1] || INT64_MAX / t.ne[3] <= t.ne[0] * t.ne[1] * t.ne[2]) {
            set_error("total number of elements in tensor '%s' is not representable", t.name.c_str()); return false;
        }
        if (!r.rd(t.type)) { set_error("failed to read tensor info"); return false; }
        if (t.type < 0 || t.type >= 40) { set_error("tensor '%s' has invalid ggml type %d", t.name.c_str(), t.type); return false; }
        const tinfo & ti = g_types[t.type];
        if (ti.qk == 0 || t.ne[0] % ti.qk != 0) {
            set_error("tensor '%s' of type %d (%s) has %lld elements per row, not a multiple of block size (%d)", t.name.c_str(), t.type, ti.name, (long long) t.ne[0], ti.qk);
            return false;
        }
        // ggml_nbytes of a contiguous tensor (src/ggml.c:1196-1219)
        t.nbytes = (uint64_t) (t.ne[0] / ti.qk) * (uint64_t) ti.bytes * (uint64_t) t.ne[1] * (uint64_t) t.ne[2] * (uint64_t) t.ne[3];
        if (!r.rd(t.offset)) { set_error("failed to read tensor info"); return false; }
        g.tensors.push_back(std::move(t));
    }
    const long pos = ftell(f);
    if (pos < 0) { set_error("failed to seek to beginning of data section"); return false; }
    g.data_offset = ((uint64_t) pos + g.alignment - 1) & ~(uint64_t) (g.alignment - 1);
    uint64_t expect = 0, dev = 0;
    for (tensor_entry & t : g.tensors) {
        if (t.offset != expect) { set_error("tensor '%s' has offset %llu, expected %llu", t.name.c_str(), (unsigned long long) t.offset, (unsigned long long) expect); return false; }
        const uint64_t padded = (t.nbytes + g.alignment - 1) & ~(uint64_t) (g.alignment - 1);
        if (UINT64_MAX - expect < padded) { set_error("tensor '%s' size overflow", t.name.c_str()); return false; }
        expect += padded;
        t.dev_offset = dev;
        dev += (t.nbytes + DEV_ALIGN - 1) & ~(uint64_t) (DEV_ALIGN - 1);
    }
    g.data_bytes = expect;
    g.device_bytes = dev;
    return true;
}

bool kv_ok(const b200_gguf * g, int64_t i) { return g && i >= 0 && i < (int64_t) g->kv.size(); }

// file -> staging -> device for a list of (file offset, bytes, device address) ranges, in list order.
struct range { uint64_t file_off; uint64_t n; char * dst; };

int load_ranges(b200_gguf * g, const std::vector<range> & ranges, cudaStream_t stream) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return -(int) e;
    for (int s = 0; s < N_STAGES; s++) {
        if (!g->stage[s]) {
            void * p = nullptr;
            if ((e = cudaHostAlloc(&p, STAGE_BYTES, cudaHostAllocPortable)) != cudaSuccess) return -(int) e;
            g->stage[s] = (char *) p;
            if ((e = cudaEventCreateWithFlags(&g->ev[s], cudaEventDisableTiming)) != cudaSuccess) return -(int) e;
        }
    }
    // Jobs: each fills one stage with up to STAGE_BYTES of consecutive pieces (a large tensor spans several jobs; small
    // tensors share one).  Pieces start on 64-byte boundaries of the stage.
    struct job { std::vector<piece> pieces; std::vector<uint64_t> file_off; };
    std::vector<job> jobs;
    {
        job cur; size_t used = 0;
        for (const range & r : ranges) {
            uint64_t done = 0;
            while (done < r.n) {
                if (used + 64 > STAGE_BYTES) { jobs.push_back(std::move(cur)); cur = job(); used = 0; }
                const size_t n = (size_t) std::min<uint64_t>(r.n - done, STAGE_BYTES - used);
                cur.pieces.push_back({used, r.dst + done, n});
                cur.file_off.push_back(r.file_off + done);
                done += n;
                used = (used + n + 63) & ~(size_t) 63;
            }
        }
        if (!cur.pieces.empty()) jobs.push_back(std::move(cur));
    }
    // stage states: 0 free, 1 filled (by the reader), then the issuing thread sends it and records ev[s]
    std::mutex mu; std::condition_variable cv;
    int state[N_STAGES] = {};
    bool read_failed = false, abort_load = false;
    auto read_stage = [&](int s) {       // reader thread of stage s: jobs s, s + N_STAGES, ...
        cudaSetDevice(dev);
        for (size_t j = (size_t) s; j < jobs.size(); j += N_STAGES) {
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return state[s] == 0 || abort_load; });
                if (abort_load) return;
            }
            if (g->ev_pending[s]) { cudaEventSynchronize(g->ev[s]); g->ev_pending[s] = false; }   // the copies that last used this stage
            bool ok = true;
            for (size_t p = 0; ok && p < jobs[j].pieces.size(); p++) {
                const piece & pc = jobs[j].pieces[p];
                size_t got = 0;
                while (got < pc.n) {
                    const ssize_t k = pread(g->fd, g->stage[s] + pc.stage_off + got, pc.n - got, (off_t) (jobs[j].file_off[p] + got));
                    if (k <= 0) { ok = false; break; }
                    got += (size_t) k;
                }
            }
            { std::lock_guard<std::mutex> lk(mu); if (ok) state[s] = 1; else read_failed = true; }
            cv.notify_all();
            if (!ok) return;
        }
    };
    std::vector<std::thread> readers;
    for (int s = 0; s < N_STAGES && (size_t) s < jobs.size(); s++) readers.emplace_back(read_stage, s);
    int rc = B200_OK;
    for (size_t j = 0; j < jobs.size() && rc == B200_OK; j++) {
        const int s = (int) (j % N_STAGES);
        {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return state[s] == 1 || read_failed; });
            if (state[s] != 1) { rc = B200_ERR_FILE; break; }
        }
        for (const piece & pc : jobs[j].pieces) {
            e = cudaMemcpyAsync(pc.dst, g->stage[s] + pc.stage_off, pc.n, cudaMemcpyHostToDevice, stream);
            if (e != cudaSuccess) { rc = -(int) e; break; }
        }
        if (rc == B200_OK && (e = cudaEventRecord(g->ev[s], stream)) != cudaSuccess) rc = -(int) e;
        { std::lock_guard<std::mutex> lk(mu); g->ev_pending[s] = rc == B200_OK; state[s] = 0; }
        cv.notify_all();
    }
    if (rc != B200_OK) {
        { std::lock_guard<std::mutex> lk(mu); abort_load = true; }
        cv.notify_all();
        if (rc == B200_ERR_FILE) set_error("read error in '%s'", g->path.c_str());
    }
    for (std::thread & t : readers) t.join();
    // the staging buffers belong to this handle: wait until the copies have left them, so that the caller may close
    // the handle (or load again) right away; the DEVICE side stays asynchronous only up to this point
    for (int s = 0; s < N_STAGES; s++) if (g->ev_pending[s]) { cudaEventSynchronize(g->ev[s]); g->ev_pending[s] = false; }
    return rc;
}

} // namespace

extern "C" const char * b200_gguf_last_error(void) { return t_error.c_str(); }

extern "C" int b200_gguf_open(const char * path, b200_gguf ** out) {
    if (!path || !out) return B200_ERR_SHAPE;
    *out = nullptr;
    FILE * f = fopen(path, "rb");
    if (!f) { set_error("failed to open GGUF file '%s'", path); return B200_ERR_FILE; }
    b200_gguf * g = new b200_gguf;
    g->path = path;
    struct stat st;
    if (fstat(fileno(f), &st) == 0) g->file_bytes = (uint64_t) st.st_size;
    const bool ok = parse(*g, f);
    fclose(f);
    if (ok && g->data_offset + g->data_bytes > g->file_bytes && !g->tensors.empty()) {
        // the reference finds this only when it reads the data (gguf.cpp:700-708: "failed to read tensor data binary blob")
        const tensor_entry & last = g->tensors.back();
        if (g->data_offset + last.offset + last.nbytes > g->file_bytes) { set_error("failed to read tensor data binary blob: file ends inside tensor '%s'", last.name.c_str()); delete g; return B200_ERR_FILE; }
    }
    if (!ok) { delete g; return B200_ERR_FILE; }
    g->fd = open(path, O_RDONLY);
    if (g->fd < 0) { set_error("failed to open GGUF file '%s'", path); delete g; return B200_ERR_FILE; }
    *out = g;
    return B200_OK;
}

extern "C" void b200_gguf_close(b200_gguf * g) {
    if (!g) return;
    for (int s = 0; s < N_STAGES; s++) {
        if (g->stage[s]) { if (g->ev_pending[s]) cudaEventSynchronize(g->ev[s]); cudaFreeHost(g->stage[s]); cudaEventDestroy(g->ev[s]); }
    }
    if (g->fd >= 0) close(g->fd);
    delete g;
}

extern "C" uint32_t b200_gguf_version(const b200_gguf * g)     { return g ? g->version : 0; }
extern "C" size_t   b200_gguf_alignment(const b200_gguf * g)   { return g ? g->alignment : 0; }
extern "C" uint64_t b200_gguf_data_offset(const b200_gguf * g) { return g ? g->data_offset : 0; }
extern "C" uint64_t b200_gguf_data_bytes(const b200_gguf * g)  { return g ? g->data_bytes : 0; }

extern "C" int64_t b200_gguf_n_kv(const b200_gguf * g) { return g ? (int64_t) g->kv.size() : 0; }
extern "C" int64_t b200_gguf_find_key(const b200_gguf * g, const char * key) {
    if (!g || !key) return -1;
    for (size_t i = 0; i < g->kv.size(); i++) if (g->kv[i].key == key) return (int64_t) i;
    return -1;
}
extern "C" const char * b200_gguf_key(const b200_gguf * g, int64_t i) { return kv_ok(g, i) ? g->kv[i].key.c_str() : nullptr; }
extern "C" int b200_gguf_kv_type(const b200_gguf * g, int64_t i)  { return kv_ok(g, i) ? (g->kv[i].is_array ? (int) KV_ARR : g->kv[i].type) : -1; }
extern "C" int b200_gguf_arr_type(const b200_gguf * g, int64_t i) { return kv_ok(g, i) && g->kv[i].is_array ? g->kv[i].type : -1; }
extern "C" uint64_t b200_gguf_arr_n(const b200_gguf * g, int64_t i) { return kv_ok(g, i) ? g->kv[i].n : 0; }
// raw little-endian bytes of a scalar or of a fixed-size-element array (NULL for strings)
extern "C" const void * b200_gguf_val_data(const b200_gguf * g, int64_t i) {
    return kv_ok(g, i) && g->kv[i].type != KV_STR ? (const void *) g->kv[i].data.data() : nullptr;
}
// j-th string of a string array; j = 0 of a scalar string
extern "C" const char * b200_gguf_val_str(const b200_gguf * g, int64_t i, uint64_t j) {
    return kv_ok(g, i) && g->kv[i].type == KV_STR && j < g->kv[i].strs.size() ? g->kv[i].strs[(size_t) j].c_str() : nullptr;
}
// any scalar numeric / bool value widened to double and int64 (the typed getters of include/gguf.h:96-108 abort on a type
// mismatch; this one returns B200_ERR_TYPE for strings and arrays)
extern "C" int b200_gguf_val_number(const b200_gguf * g, int64_t i, double * as_f64, int64_t * as_i64) {
    if (!kv_ok(g, i)) return B200_ERR_SHAPE;
    const kv_entry & e = g->kv[i];
    if (e.is_array || e.type == KV_STR) return B200_ERR_TYPE;
    const uint8_t * p = e.data.data();
    double d = 0; int64_t v = 0;
    switch (e.type) {
        case KV_U8:  { uint8_t x;  memcpy(&x, p, 1); v = x; d = x; } break;
        case KV_I8:  { int8_t x;   memcpy(&x, p, 1); v = x; d = x; } break;
        case KV_BOOL:{ int8_t x;   memcpy(&x, p, 1); v = x != 0; d = (double) v; } break;
        case KV_U16: { uint16_t x; memcpy(&x, p, 2); v = x; d = x; } break;
        case KV_I16: { int16_t x;  memcpy(&x, p, 2); v = x; d = x; } break;
        case KV_U32: { uint32_t x; memcpy(&x, p, 4); v = x; d = x; } break;
        case KV_I32: { int32_t x;  memcpy(&x, p, 4); v = x; d = x; } break;
        case KV_F32: { float x;    memcpy(&x, p, 4); v = (int64_t) x; d = x; } break;
        case KV_U64: { uint64_t x; memcpy(&x, p, 8); v = (int64_t) x; d = (double) x; } break;
        case KV_I64: { int64_t x;  memcpy(&x, p, 8); v = x; d = (double) x; } break;
        case KV_F64: { double x;   memcpy(&x, p, 8); v = (int64_t) x; d = x; } break;
        default: return B200_ERR_TYPE;
    }
    if (as_f64) *as_f64 = d;
    if (as_i64) *as_i64 = v;
    return B200_OK;
}

extern "C" int64_t b200_gguf_n_tensors(const b200_gguf * g) { return g ? (int64_t) g->tensors.size() : 0; }
extern "C" int64_t b200_gguf_find_tensor(const b200_gguf * g, const char * name) {
    if (!g || !name) return -1;
    for (size_t i = 0; i < g->tensors.size(); i++) if (g->tensors[i].name == name) return (int64_t) i;
    return -1;
}
extern "C" int b200_gguf_tensor(const b200_gguf * g, int64_t i, b200_gguf_tensor_info * out) {
    if (!g || !out || i < 0 || i >= (int64_t) g->tensors.size()) return B200_ERR_SHAPE;
    const tensor_entry & t = g->tensors[(size_t) i];
    out->name = t.name.c_str(); out->type = t.type; out->n_dims = t.n_dims;
    for (int d = 0; d < 4; d++) out->ne[d] = t.ne[d];
    out->file_offset = g->data_offset + t.offset; out->nbytes = t.nbytes; out->device_offset = t.dev_offset;
    return B200_OK;
}
extern "C" size_t b200_gguf_device_bytes(const b200_gguf * g) { return g ? (size_t) g->device_bytes : 0; }
extern "C" size_t b200_gguf_type_row_bytes(int type, int64_t ne0) {
    if (type < 0 || type >= 40 || g_types[type].qk == 0 || ne0 < 0 || ne0 % g_types[type].qk != 0) return 0;
    return (size_t) (ne0 / g_types[type].qk) * (size_t) g_types[type].bytes;
}

extern "C" int b200_gguf_load(b200_gguf * g, void * base_dev, size_t bytes, void * stream) {
    if (!g || (!base_dev && g->device_bytes)) return B200_ERR_SHAPE;
    if (bytes < g->device_bytes) return B200_ERR_WORKSPACE;
    if ((uintptr_t) base_dev % DEV_ALIGN != 0) return B200_ERR_STRIDE;
    std::vector<range> ranges;
    ranges.reserve(g->tensors.size());
    for (const tensor_entry & t : g->tensors)
        if (t.nbytes) ranges.push_back({g->data_offset + t.offset, t.nbytes, (char *) base_dev + t.dev_offset});
    return load_ranges(g, ranges, (cudaStream_t) stream);
}

extern "C" int b200_gguf_load_tensors(b200_gguf * g, const int64_t * ids, void * const * dst_dev, int64_t count, void * stream) {
    if (!g || count < 0 || (count && (!ids || !dst_dev))) return B200_ERR_SHAPE;
    std::vector<range> ranges;
    ranges.reserve((size_t) count);
    for (int64_t c = 0; c < count; c++) {
        if (ids[c] < 0 || ids[c] >= (int64_t) g->tensors.size() || !dst_dev[c]) return B200_ERR_SHAPE;
        const tensor_entry & t = g->tensors[(size_t) ids[c]];
        if (t.nbytes) ranges.push_back({g->data_offset + t.offset, t.nbytes, (char *) dst_dev[c]});
    }
    return load_ranges(g, ranges, (cudaStream_t) stream);
}
