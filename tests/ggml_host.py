"""TEST / BENCH HARNESS (not product code).  ctypes view of ggml's PUBLIC C API (include/ggml.h, include/ggml-backend.h, include/ggml-alloc.h).

This is the *host application* side of the drop-in boundary: the same calls an unmodified ggml
program makes (examples/simple/simple-backend.cpp:50-140 — init context, build a graph with
ggml_mul_mat, ggml_backend_alloc_ctx_tensors, ggml_backend_tensor_set, ggml_backend_graph_compute,
ggml_backend_tensor_get).  It is pointed at a directory holding a ggml build (libggml-base.so,
libggml.so, libggml-cpu-*.so); here that is oracle/_ref/bin, the unmodified reference compiled by
oracle/Makefile.  Our plugin (libggml-b200.so) is loaded into it with ggml_backend_load(), which is
what GGML_BACKEND_PATH does (src/ggml-backend-reg.cpp:595-599).

Nothing in here computes anything: it only marshals calls into ggml.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

# enum ggml_type (include/ggml.h:370-412)
GGML_TYPE = {
    "f32": 0, "f16": 1, "q4_0": 2, "q4_1": 3, "q5_0": 6, "q5_1": 7, "q8_0": 8, "q8_1": 9,
    "q2_K": 10, "q3_K": 11, "q4_K": 12, "q5_K": 13, "q6_K": 14, "q8_K": 15, "iq4_nl": 20, "i32": 26,
}
BLOCK_ELEMS = {"f32": 1, "f16": 1, "i32": 1, "q4_0": 32, "q8_0": 32, "q8_1": 32, "q4_K": 256, "q5_K": 256, "q6_K": 256,
               "q4_1": 32, "q5_0": 32, "q5_1": 32, "iq4_nl": 32}
BLOCK_BYTES = {"f32": 4, "f16": 2, "i32": 4, "q4_0": 18, "q8_0": 34, "q8_1": 36, "q4_K": 144, "q5_K": 176, "q6_K": 210,
               "q4_1": 20, "q5_0": 22, "q5_1": 24, "iq4_nl": 18}
QUANT_TYPES = ("q4_0", "q8_0", "q4_K", "q5_K", "q6_K")
EXTRA_TYPES = ("q4_1", "q5_0", "q5_1", "iq4_nl")

GGML_BACKEND_DEVICE_TYPE_CPU = 0
GGML_BACKEND_DEVICE_TYPE_GPU = 1
GGML_STATUS_SUCCESS = 0


def row_bytes(tname: str, k: int) -> int:
    assert k % BLOCK_ELEMS[tname] == 0, (tname, k)
    return k // BLOCK_ELEMS[tname] * BLOCK_BYTES[tname]


class _InitParams(C.Structure):
    _fields_ = [("mem_size", C.c_size_t), ("mem_buffer", C.c_void_p), ("no_alloc", C.c_bool)]


def default_ggml_dir() -> str:
    here = os.path.dirname(os.path.abspath(__file__))
    return os.environ.get("GGML_B200_HOST_DIR", os.path.join(here, "..", "oracle", "_ref", "bin"))   # tests/ -> repo root


class GgmlHost:
    """One loaded ggml (base + registry).  Backends are loaded lazily."""

    def __init__(self, ggml_dir: Optional[str] = None):
        self.dir = os.path.abspath(ggml_dir or default_ggml_dir())
        base = os.path.join(self.dir, "libggml-base.so")
        if not os.path.exists(base):
            raise FileNotFoundError(f"no ggml build at {self.dir} (run `make -C oracle ref` where /root/reference exists)")
        self.base = C.CDLL(base, mode=C.RTLD_GLOBAL)
        self.reg = C.CDLL(os.path.join(self.dir, "libggml.so"), mode=C.RTLD_GLOBAL)
        self._cpu_loaded = False
        self._proto()

    # ------------------------------------------------------------------ prototypes
    def _proto(self):
        b, r = self.base, self.reg
        vp, i64, sz = C.c_void_p, C.c_int64, C.c_size_t
        b.ggml_init.restype = vp;                       b.ggml_init.argtypes = [_InitParams]
        b.ggml_free.restype = None;                     b.ggml_free.argtypes = [vp]
        b.ggml_tensor_overhead.restype = sz
        b.ggml_graph_overhead.restype = sz
        b.ggml_new_tensor_1d.restype = vp;              b.ggml_new_tensor_1d.argtypes = [vp, C.c_int, i64]
        b.ggml_new_tensor_2d.restype = vp;              b.ggml_new_tensor_2d.argtypes = [vp, C.c_int, i64, i64]
        b.ggml_new_tensor_4d.restype = vp;              b.ggml_new_tensor_4d.argtypes = [vp, C.c_int, i64, i64, i64, i64]
        b.ggml_mul_mat.restype = vp;                    b.ggml_mul_mat.argtypes = [vp, vp, vp]
        b.ggml_get_rows.restype = vp;                   b.ggml_get_rows.argtypes = [vp, vp, vp]
        b.ggml_cpy.restype = vp;                        b.ggml_cpy.argtypes = [vp, vp, vp]
        b.ggml_new_graph.restype = vp;                  b.ggml_new_graph.argtypes = [vp]
        b.ggml_new_graph_custom.restype = vp;           b.ggml_new_graph_custom.argtypes = [vp, sz, C.c_bool]
        b.ggml_build_forward_expand.restype = None;     b.ggml_build_forward_expand.argtypes = [vp, vp]
        b.ggml_nbytes.restype = sz;                     b.ggml_nbytes.argtypes = [vp]
        b.ggml_nelements.restype = i64;                 b.ggml_nelements.argtypes = [vp]
        b.ggml_quantize_chunk.restype = sz
        b.ggml_quantize_chunk.argtypes = [C.c_int, vp, vp, i64, i64, i64, vp]
        b.ggml_fp16_to_fp32_row.restype = None;         b.ggml_fp16_to_fp32_row.argtypes = [vp, vp, i64]
        b.ggml_fp32_to_fp16_row.restype = None;         b.ggml_fp32_to_fp16_row.argtypes = [vp, vp, i64]
        b.quantize_row_q8_1_ref.restype = None;         b.quantize_row_q8_1_ref.argtypes = [vp, vp, i64]
        b.quantize_row_q8_K_ref.restype = None;         b.quantize_row_q8_K_ref.argtypes = [vp, vp, i64]
        for t in QUANT_TYPES + EXTRA_TYPES:
            f = getattr(b, f"dequantize_row_{t}")
            f.restype = None; f.argtypes = [vp, vp, i64]
        b.ggml_backend_alloc_ctx_tensors.restype = vp;  b.ggml_backend_alloc_ctx_tensors.argtypes = [vp, vp]
        b.ggml_backend_alloc_ctx_tensors_from_buft.restype = vp
        b.ggml_backend_alloc_ctx_tensors_from_buft.argtypes = [vp, vp]
        b.ggml_backend_buffer_free.restype = None;      b.ggml_backend_buffer_free.argtypes = [vp]
        b.ggml_backend_tensor_set.restype = None;       b.ggml_backend_tensor_set.argtypes = [vp, vp, sz, sz]
        b.ggml_backend_tensor_get.restype = None;       b.ggml_backend_tensor_get.argtypes = [vp, vp, sz, sz]
        b.ggml_backend_graph_compute.restype = C.c_int; b.ggml_backend_graph_compute.argtypes = [vp, vp]
        b.ggml_backend_graph_compute_async.restype = C.c_int
        b.ggml_backend_graph_compute_async.argtypes = [vp, vp]
        b.ggml_backend_synchronize.restype = None;      b.ggml_backend_synchronize.argtypes = [vp]
        b.ggml_backend_free.restype = None;             b.ggml_backend_free.argtypes = [vp]
        b.ggml_backend_name.restype = C.c_char_p;       b.ggml_backend_name.argtypes = [vp]
        b.ggml_backend_dev_init.restype = vp;           b.ggml_backend_dev_init.argtypes = [vp, C.c_char_p]
        b.ggml_backend_dev_name.restype = C.c_char_p;   b.ggml_backend_dev_name.argtypes = [vp]
        b.ggml_backend_dev_description.restype = C.c_char_p; b.ggml_backend_dev_description.argtypes = [vp]
        b.ggml_backend_dev_supports_op.restype = C.c_bool;   b.ggml_backend_dev_supports_op.argtypes = [vp, vp]
        b.ggml_backend_dev_buffer_type.restype = vp;    b.ggml_backend_dev_buffer_type.argtypes = [vp]
        b.ggml_backend_get_device.restype = vp;         b.ggml_backend_get_device.argtypes = [vp]
        b.ggml_backend_dev_backend_reg.restype = vp;    b.ggml_backend_dev_backend_reg.argtypes = [vp]
        b.ggml_backend_reg_get_proc_address.restype = vp
        b.ggml_backend_reg_get_proc_address.argtypes = [vp, C.c_char_p]
        b.ggml_backend_reg_name.restype = C.c_char_p;   b.ggml_backend_reg_name.argtypes = [vp]
        b.ggml_backend_reg_dev_count.restype = sz;      b.ggml_backend_reg_dev_count.argtypes = [vp]
        b.ggml_backend_reg_dev_get.restype = vp;        b.ggml_backend_reg_dev_get.argtypes = [vp, sz]
        r.ggml_backend_load.restype = vp;               r.ggml_backend_load.argtypes = [C.c_char_p]
        r.ggml_backend_load_all_from_path.restype = None
        r.ggml_backend_load_all_from_path.argtypes = [C.c_char_p]
        r.ggml_backend_dev_count.restype = sz
        r.ggml_backend_dev_get.restype = vp;            r.ggml_backend_dev_get.argtypes = [sz]
        r.ggml_backend_dev_by_name.restype = vp;        r.ggml_backend_dev_by_name.argtypes = [C.c_char_p]
        r.ggml_backend_init_by_type.restype = vp;       r.ggml_backend_init_by_type.argtypes = [C.c_int, C.c_char_p]
        r.ggml_backend_init_by_name.restype = vp;       r.ggml_backend_init_by_name.argtypes = [C.c_char_p, C.c_char_p]

    # ------------------------------------------------------------------ backends
    def load_cpu(self):
        if not self._cpu_loaded:
            # the reference's own variant picker (ggml_backend_load_best("cpu", ..)) over libggml-cpu-*.so
            self.reg.ggml_backend_load_all_from_path(self.dir.encode())
            self._cpu_loaded = True

    def cpu_backend(self, n_threads: Optional[int] = None):
        self.load_cpu()
        be = self.reg.ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, None)
        if not be:
            raise RuntimeError("ggml CPU backend failed to load")
        if n_threads:
            dev = self.base.ggml_backend_get_device(be)
            reg = self.base.ggml_backend_dev_backend_reg(dev)
            fn = self.base.ggml_backend_reg_get_proc_address(reg, b"ggml_backend_set_n_threads")
            if fn:
                C.CFUNCTYPE(None, C.c_void_p, C.c_int)(fn)(be, int(n_threads))
        return be

    def load_plugin(self, path: str):
        """ggml_backend_load(path): what GGML_BACKEND_PATH triggers (src/ggml-backend-reg.cpp:595-599)."""
        reg = self.reg.ggml_backend_load(os.path.abspath(path).encode())
        if not reg:
            raise RuntimeError(f"ggml_backend_load({path}) failed")
        return reg

    def device_by_name(self, name: str):
        return self.reg.ggml_backend_dev_by_name(name.encode())

    def init_device(self, dev):
        be = self.base.ggml_backend_dev_init(dev, None)
        if not be:
            raise RuntimeError("ggml_backend_dev_init failed")
        return be

    # ------------------------------------------------------------------ quantization helpers (ggml-base)
    def quantize(self, tname: str, x: np.ndarray, imatrix: Optional[np.ndarray] = None) -> np.ndarray:
        """ggml_quantize_chunk on rows of x (2-D f32) -> uint8 [nrows, row_bytes]."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        nrows, k = x.shape
        out = np.empty((nrows, row_bytes(tname, k)), dtype=np.uint8)
        im = None
        if imatrix is not None:
            imatrix = np.ascontiguousarray(imatrix, dtype=np.float32)
            im = imatrix.ctypes.data
        n = self.base.ggml_quantize_chunk(GGML_TYPE[tname], x.ctypes.data, out.ctypes.data, 0, nrows, k, im)
        assert n == out.nbytes, (n, out.nbytes)
        return out

    def dequantize(self, tname: str, blocks: np.ndarray, k_total: int) -> np.ndarray:
        blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
        out = np.empty(k_total, dtype=np.float32)
        getattr(self.base, f"dequantize_row_{tname}")(blocks.ctypes.data, out.ctypes.data, k_total)
        return out

    def quantize_q8_1(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
        out = np.empty(x.size // 32 * 36, dtype=np.uint8)
        self.base.quantize_row_q8_1_ref(x.ctypes.data, out.ctypes.data, x.size)
        return out

    def quantize_q8_K(self, x: np.ndarray):
        """quantize_row_q8_K_ref -> (q int8 [k], d f32 [k/256], bsums int16 [k/16]) taken out of the block_q8_K structs
        (f32 d | i8 qs[256] | i16 bsums[16] = 292 bytes, src/ggml-common.h:338-343)"""
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
        nb = x.size // 256
        out = np.zeros(nb * 292, dtype=np.uint8)
        self.base.quantize_row_q8_K_ref(x.ctypes.data, out.ctypes.data, x.size)
        blk = out.reshape(nb, 292)
        d = np.ascontiguousarray(blk[:, 0:4]).view(np.float32).reshape(-1)
        q = np.ascontiguousarray(blk[:, 4:260]).view(np.int8).reshape(-1)
        bs = np.ascontiguousarray(blk[:, 260:292]).view(np.int16).reshape(-1)
        return q, d, bs

    def fp16_to_fp32(self, h: np.ndarray) -> np.ndarray:
        h = np.ascontiguousarray(h, dtype=np.uint16)
        out = np.empty(h.size, dtype=np.float32)
        self.base.ggml_fp16_to_fp32_row(h.ctypes.data, out.ctypes.data, h.size)
        return out

    def fp32_to_fp16(self, f: np.ndarray) -> np.ndarray:
        f = np.ascontiguousarray(f, dtype=np.float32)
        out = np.empty(f.size, dtype=np.uint16)
        self.base.ggml_fp32_to_fp16_row(f.ctypes.data, out.ctypes.data, f.size)
        return out

    # ------------------------------------------------------------------ graphs
    def new_context(self, n_tensors: int = 64, graph_nodes: int = 0):
        size = self.base.ggml_tensor_overhead() * (n_tensors + 8) + self.base.ggml_graph_overhead() * 2
        if graph_nodes:
            size += graph_nodes * 64 + self.base.ggml_graph_overhead()
        ctx = self.base.ggml_init(_InitParams(size, None, True))
        if not ctx:
            raise MemoryError("ggml_init")
        return ctx

    def tensor_set(self, t, arr: np.ndarray, offset: int = 0):
        arr = np.ascontiguousarray(arr)
        self.base.ggml_backend_tensor_set(t, arr.ctypes.data, offset, arr.nbytes)

    def tensor_get(self, t, dtype=np.float32) -> np.ndarray:
        n = self.base.ggml_nbytes(t)
        out = np.empty(n // np.dtype(dtype).itemsize, dtype=dtype)
        self.base.ggml_backend_tensor_get(t, out.ctypes.data, 0, n)
        return out

    def mul_mat(self, backend, tname: str, w_blocks: np.ndarray, m: int, k: int, x: np.ndarray) -> np.ndarray:
        """dst[n, m] = ggml_mul_mat(W[k, m] as `tname`, X[k, n] f32) on `backend`; returns [n, m] f32."""
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, k)
        n = x.shape[0]
        b = self.base
        ctx = self.new_context()
        try:
            a = b.ggml_new_tensor_2d(ctx, GGML_TYPE[tname], k, m)
            bb = b.ggml_new_tensor_2d(ctx, GGML_TYPE["f32"], k, n)
            c = b.ggml_mul_mat(ctx, a, bb)
            gf = b.ggml_new_graph(ctx)
            b.ggml_build_forward_expand(gf, c)
            buf = b.ggml_backend_alloc_ctx_tensors(ctx, backend)
            if not buf:
                raise MemoryError("ggml_backend_alloc_ctx_tensors")
            try:
                self.tensor_set(a, np.ascontiguousarray(w_blocks, dtype=np.uint8))
                self.tensor_set(bb, x)
                st = b.ggml_backend_graph_compute(backend, gf)
                if st != GGML_STATUS_SUCCESS:
                    raise RuntimeError(f"graph_compute status {st}")
                return self.tensor_get(c).reshape(n, m)
            finally:
                b.ggml_backend_buffer_free(buf)
        finally:
            b.ggml_free(ctx)


def split_buffer_type(host: "GgmlHost", backend, main_device: int, tensor_split: Sequence[float]):
    """The backend registry's "ggml_backend_split_buffer_type" proc (include/ggml-backend.h:188), as llama.cpp
    obtains it for --split-mode row."""
    b = host.base
    reg = b.ggml_backend_dev_backend_reg(b.ggml_backend_get_device(backend))
    fn = b.ggml_backend_reg_get_proc_address(reg, b"ggml_backend_split_buffer_type")
    if not fn:
        return None
    ts = (C.c_float * 16)(*([float(v) for v in tensor_split] + [0.0] * (16 - len(tensor_split))))
    return C.CFUNCTYPE(C.c_void_p, C.c_int, C.POINTER(C.c_float))(fn)(int(main_device), ts)


def mul_mat_split(host: "GgmlHost", backend, buft, tname: str, w_blocks: np.ndarray, m: int, k: int, x: np.ndarray) -> np.ndarray:
    """ggml_mul_mat with the weight tensor allocated in the row-split buffer type `buft` and the activations /
    output on `backend` (the main device): returns [n, m] f32, and the weight bytes read back through get_tensor."""
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, k)
    n = x.shape[0]
    b = host.base
    ctx_w, ctx = host.new_context(), host.new_context()
    buf_w = buf = None
    try:
        a = b.ggml_new_tensor_2d(ctx_w, GGML_TYPE[tname], k, m)
        buf_w = b.ggml_backend_alloc_ctx_tensors_from_buft(ctx_w, buft)
        if not buf_w:
            raise MemoryError("ggml_backend_alloc_ctx_tensors_from_buft(split)")
        bb = b.ggml_new_tensor_2d(ctx, GGML_TYPE["f32"], k, n)
        c = b.ggml_mul_mat(ctx, a, bb)
        gf = b.ggml_new_graph(ctx)
        b.ggml_build_forward_expand(gf, c)
        buf = b.ggml_backend_alloc_ctx_tensors(ctx, backend)
        if not buf:
            raise MemoryError("ggml_backend_alloc_ctx_tensors")
        host.tensor_set(a, np.ascontiguousarray(w_blocks, dtype=np.uint8))
        host.tensor_set(bb, x)
        st = b.ggml_backend_graph_compute(backend, gf)
        if st != GGML_STATUS_SUCCESS:
            raise RuntimeError(f"graph_compute status {st}")
        return host.tensor_get(c).reshape(n, m), host.tensor_get(a, dtype=np.uint8)
    finally:
        if buf:
            b.ggml_backend_buffer_free(buf)
        if buf_w:
            b.ggml_backend_buffer_free(buf_w)
        b.ggml_free(ctx); b.ggml_free(ctx_w)


class MulMatGraph:
    """A ggml graph of independent MUL_MAT nodes (one per (type, m, k) entry, f32 activations [k, n])
    allocated once on `backend`: weights / activations are set with ggml_backend_tensor_set, compute()
    is ggml_backend_graph_compute — the calls an unmodified ggml program makes per step."""

    def __init__(self, host: GgmlHost, backend, mats: Sequence, n: int):
        self.host, self.backend, self.mats, self.n = host, backend, list(mats), n
        b = host.base
        self.ctx = host.new_context(n_tensors=3 * len(self.mats) + 8, graph_nodes=2 * len(self.mats) + 64)
        self.w, self.x, self.y = [], [], []
        for mt in self.mats:
            a = b.ggml_new_tensor_2d(self.ctx, GGML_TYPE[mt.type], mt.k, mt.m)
            x = b.ggml_new_tensor_2d(self.ctx, GGML_TYPE["f32"], mt.k, n)
            self.w.append(a); self.x.append(x); self.y.append(b.ggml_mul_mat(self.ctx, a, x))
        self.graph = b.ggml_new_graph_custom(self.ctx, max(2048, 2 * len(self.mats) + 64), False)
        for y in self.y:
            b.ggml_build_forward_expand(self.graph, y)
        self.buf = b.ggml_backend_alloc_ctx_tensors(self.ctx, backend)
        if not self.buf:
            b.ggml_free(self.ctx)
            raise MemoryError("ggml_backend_alloc_ctx_tensors")

    def set_weight(self, i: int, blocks: np.ndarray):
        self.host.tensor_set(self.w[i], np.ascontiguousarray(blocks, dtype=np.uint8))

    def set_input(self, i: int, x: np.ndarray):
        self.host.tensor_set(self.x[i], np.ascontiguousarray(x, dtype=np.float32))

    def get_output(self, i: int) -> np.ndarray:
        return self.host.tensor_get(self.y[i]).reshape(self.n, self.mats[i].m)

    def compute(self):
        st = self.host.base.ggml_backend_graph_compute(self.backend, self.graph)
        if st != GGML_STATUS_SUCCESS:
            raise RuntimeError(f"graph_compute status {st}")

    def close(self):
        if self.buf:
            self.host.base.ggml_backend_buffer_free(self.buf)
            self.host.base.ggml_free(self.ctx)
            self.buf = None
