"""ctypes binding of libggml_b200_kernels.so (the C ABI in include/ggml_b200.h).

This is the host-side mirror of the reference's kernel launchers for the quantized hot path
(src/ggml-cuda/{convert,quantize,getrows,mmvq,mmq}.cu host functions): same operations, same
argument meaning (ggml ne/nb conventions), errors raised instead of GGML_ABORT.  PyTorch is used
only as plumbing: device buffers (uint8 / float32 tensors), streams and torch.distributed.  Nothing
here computes on the CPU; if the library or a device is missing the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GGML_B200_KERNELS_LIB") or os.path.join(HERE, "libggml_b200_kernels.so")   # override: tuning builds

TYPE_ID = {"f32": 0, "f16": 1, "q4_0": 2, "q8_0": 8, "q8_1": 9, "q4_K": 12, "q5_K": 13, "q6_K": 14, "i32": 26, "bf16": 30,
           "q4_1": 3, "q5_0": 6, "q5_1": 7, "iq4_nl": 20}
BLOCK_ELEMS = {"q4_0": 32, "q8_0": 32, "q8_1": 32, "q4_K": 256, "q5_K": 256, "q6_K": 256, "q4_1": 32, "q5_0": 32, "q5_1": 32, "iq4_nl": 32}
BLOCK_BYTES = {"q4_0": 18, "q8_0": 34, "q8_1": 36, "q4_K": 144, "q5_K": 176, "q6_K": 210, "q4_1": 20, "q5_0": 22, "q5_1": 24, "iq4_nl": 18}
QUANT_TYPES = ("q4_0", "q8_0", "q4_K", "q5_K", "q6_K")           # the hot path: every kernel
EXTRA_TYPES = ("q4_1", "q5_0", "q5_1", "iq4_nl")                 # dequantize / GET_ROWS / CPY and the general mat-vec only
MMVQ_MAX_N = 8
# enum b200_option (include/ggml_b200.h)
OPT = {"mmvq_generic": 0, "mmvq_no_fused": 1, "batch_single": 2, "no_batch_kernel": 3, "mmq_bn": 4,
       "mmq_force_unaligned": 5, "no_chain": 6, "mmq_route": 7}
ERR_FLAG_INDEX = 1


class B200Error(RuntimeError):
    pass


def row_bytes(t: str, k: int) -> int:
    if k % BLOCK_ELEMS[t] != 0:
        raise ValueError(f"k={k} is not a multiple of the {t} block size")
    return k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]


class MulMatNode(C.Structure):
    """struct b200_mul_mat_node (include/ggml_b200.h)."""
    _fields_ = [("type_a", C.c_int32), ("flags", C.c_uint32), ("a_dev", C.c_void_p),
                ("ne_a", C.c_int64 * 4), ("nb_a", C.c_size_t * 4), ("b_dev", C.c_void_p),
                ("ne_b", C.c_int64 * 4), ("nb_b", C.c_size_t * 4), ("dst_dev", C.c_void_p)]


class PeerGroup(C.Structure):
    """struct b200_peer_group (include/ggml_b200.h): base[r] = this process's mapping of rank r's symmetric region."""
    _fields_ = [("n_ranks", C.c_int32), ("rank", C.c_int32), ("base", C.c_void_p * 8), ("bytes", C.c_size_t)]


class GgufTensorInfo(C.Structure):
    """struct b200_gguf_tensor_info (include/ggml_b200.h)."""
    _fields_ = [("name", C.c_char_p), ("type", C.c_int32), ("n_dims", C.c_int32), ("ne", C.c_int64 * 4),
                ("file_offset", C.c_uint64), ("nbytes", C.c_uint64), ("device_offset", C.c_uint64)]


def peer_group(rank: int, bases, nbytes: int) -> PeerGroup:
    pg = PeerGroup()
    pg.n_ranks = len(bases); pg.rank = rank; pg.bytes = nbytes
    for r, b in enumerate(bases):
        pg.base[r] = int(b)
    return pg


def mul_mat_node(t: str, a_ptr: int, m: int, k: int, b_ptr: int, n: int, dst_ptr: int,
                 a_row_stride: Optional[int] = None, b_row_stride: Optional[int] = None, const_weights: bool = False) -> MulMatNode:
    """A plain 2-D node: W[m rows of k] x X[n rows of k] -> dst[n][m]."""
    rb = a_row_stride or row_bytes(t, k)
    sb = b_row_stride or 4 * k
    nd = MulMatNode()
    nd.type_a = TYPE_ID[t]; nd.flags = 1 if const_weights else 0; nd.a_dev = a_ptr; nd.b_dev = b_ptr; nd.dst_dev = dst_ptr
    nd.ne_a[:] = [k, m, 1, 1]; nd.nb_a[:] = [BLOCK_BYTES[t], rb, rb * m, rb * m]
    nd.ne_b[:] = [k, n, 1, 1]; nd.nb_b[:] = [4, sb, sb * n, sb * n]
    return nd


def load_library(path: Optional[str] = None) -> C.CDLL:
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise B200Error(f"{path} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        f"(there is no CPU fallback for these ops)")
    lib = C.CDLL(path)
    vp, i64, sz, i32 = C.c_void_p, C.c_int64, C.c_size_t, C.c_int
    A4i, A4s = C.POINTER(i64), C.POINTER(sz)
    lib.b200_abi_version.restype = i32
    lib.b200_error_string.restype = C.c_char_p;  lib.b200_error_string.argtypes = [i32]
    lib.b200_device_count.restype = i32
    lib.b200_device_info.restype = i32
    lib.b200_device_info.argtypes = [i32, C.c_char_p, sz, C.POINTER(sz), C.POINTER(sz), C.POINTER(i32), C.POINTER(i32)]
    lib.b200_block_elems.restype = i32;          lib.b200_block_elems.argtypes = [i32]
    lib.b200_block_bytes.restype = sz;           lib.b200_block_bytes.argtypes = [i32]
    lib.b200_row_bytes.restype = sz;             lib.b200_row_bytes.argtypes = [i32, i64]
    lib.b200_dequantize_row.restype = i32;       lib.b200_dequantize_row.argtypes = [i32, vp, vp, i64, i32, vp]
    lib.b200_dequantize_4d.restype = i32
    lib.b200_dequantize_4d.argtypes = [i32, vp, A4i, A4s, vp, i32, A4i, A4s, vp]
    lib.b200_quantize_row_q8_1.restype = i32;    lib.b200_quantize_row_q8_1.argtypes = [vp, sz, i64, i64, vp, vp]
    lib.b200_quantize_rows.restype = i32;        lib.b200_quantize_rows.argtypes = [i32, vp, sz, i64, i64, vp, sz, vp]
    lib.b200_quantize_rows_4d.restype = i32;     lib.b200_quantize_rows_4d.argtypes = [i32, vp, A4i, A4s, vp, A4s, vp]
    lib.b200_mul_mat_id_workspace_size.restype = sz
    lib.b200_mul_mat_id_workspace_size.argtypes = [i32, A4i, A4i]
    lib.b200_mul_mat_id.restype = i32
    lib.b200_mul_mat_id.argtypes = [i32, vp, A4i, A4s, vp, A4i, A4s, vp, C.POINTER(i64), C.POINTER(sz), vp, vp, sz, vp]
    lib.b200_get_rows.restype = i32
    lib.b200_get_rows.argtypes = [i32, vp, sz, i64, i64, vp, sz, i64, vp, i32, sz, vp]
    lib.b200_mul_mat_workspace_size.restype = sz
    lib.b200_mul_mat_workspace_size.argtypes = [i32, A4i, A4i]
    lib.b200_mul_mat.restype = i32
    lib.b200_mul_mat.argtypes = [i32, vp, A4i, A4s, vp, A4i, A4s, vp, vp, sz, vp]
    lib.b200_mul_mat_ex.restype = i32
    lib.b200_mul_mat_ex.argtypes = [i32, vp, A4i, A4s, vp, A4i, A4s, vp, vp, sz, C.c_uint32, vp]
    lib.b200_acts_scratch_bytes.restype = sz;    lib.b200_acts_scratch_bytes.argtypes = [i64, i64]
    lib.b200_quantize_acts.restype = i32;        lib.b200_quantize_acts.argtypes = [vp, sz, i64, i64, vp, vp]
    lib.b200_mul_mat_vec_q.restype = i32
    lib.b200_mul_mat_vec_q.argtypes = [i32, vp, sz, i64, i64, vp, i64, vp, sz, vp]
    lib.b200_static_scratch_bytes.restype = sz
    lib.b200_static_scratch_bytes.argtypes = []
    lib.b200_mul_mat_vec_static.restype = i32
    lib.b200_mul_mat_vec_static.argtypes = [i32, vp, sz, i64, i64, vp, sz, i64, vp, sz, vp]
    lib.b200_mul_mat_q_workspace_size.restype = sz
    lib.b200_mul_mat_q_workspace_size.argtypes = [i64, i64, i64]
    lib.b200_mul_mat_q.restype = i32
    lib.b200_mul_mat_q.argtypes = [i32, vp, sz, i64, i64, vp, sz, i64, vp, sz, vp, sz, vp]
    lib.b200_mul_mat_batch_workspace_size.restype = sz
    lib.b200_mul_mat_batch_workspace_size.argtypes = [C.POINTER(MulMatNode), i64]
    lib.b200_mul_mat_batch.restype = i32
    lib.b200_mul_mat_batch.argtypes = [C.POINTER(MulMatNode), i64, vp, sz, vp]
    lib.b200_peer_ctrl_bytes.restype = sz
    lib.b200_mul_mat_batch_peers.restype = i32
    lib.b200_mul_mat_batch_peers.argtypes = [C.POINTER(MulMatNode), i64, C.POINTER(PeerGroup), vp, sz, vp]
    lib.b200_mul_mat_batch_plan.restype = i32
    lib.b200_mul_mat_batch_plan.argtypes = [C.POINTER(MulMatNode), i64, C.POINTER(C.c_int32)]
    lib.b200_launch_count.restype = C.c_uint64
    lib.b200_acts_record_bytes.restype = sz;     lib.b200_acts_record_bytes.argtypes = [i64]
    lib.b200_acts_columns_bytes.restype = sz;    lib.b200_acts_columns_bytes.argtypes = [i64, i64]
    lib.b200_quantize_acts_record.restype = i32; lib.b200_quantize_acts_record.argtypes = [vp, i64, vp, vp]
    lib.b200_quantize_acts_columns.restype = i32; lib.b200_quantize_acts_columns.argtypes = [vp, sz, i64, i64, vp, vp]
    lib.b200_tensor_peak.restype = i32
    lib.b200_tensor_peak.argtypes = [i32, vp, vp, vp, i32, i32, C.POINTER(C.c_double), vp]
    lib.b200_set_option.restype = i32;           lib.b200_set_option.argtypes = [i32, i32]
    lib.b200_device_error_flags.restype = C.c_uint; lib.b200_device_error_flags.argtypes = [i32]
    u64 = C.c_uint64
    lib.b200_gguf_open.restype = i32;            lib.b200_gguf_open.argtypes = [C.c_char_p, C.POINTER(vp)]
    lib.b200_gguf_close.restype = None;          lib.b200_gguf_close.argtypes = [vp]
    lib.b200_gguf_last_error.restype = C.c_char_p
    lib.b200_gguf_version.restype = C.c_uint32;  lib.b200_gguf_version.argtypes = [vp]
    lib.b200_gguf_alignment.restype = sz;        lib.b200_gguf_alignment.argtypes = [vp]
    lib.b200_gguf_data_offset.restype = u64;     lib.b200_gguf_data_offset.argtypes = [vp]
    lib.b200_gguf_data_bytes.restype = u64;      lib.b200_gguf_data_bytes.argtypes = [vp]
    lib.b200_gguf_n_kv.restype = i64;            lib.b200_gguf_n_kv.argtypes = [vp]
    lib.b200_gguf_find_key.restype = i64;        lib.b200_gguf_find_key.argtypes = [vp, C.c_char_p]
    lib.b200_gguf_key.restype = C.c_char_p;      lib.b200_gguf_key.argtypes = [vp, i64]
    lib.b200_gguf_kv_type.restype = i32;         lib.b200_gguf_kv_type.argtypes = [vp, i64]
    lib.b200_gguf_arr_type.restype = i32;        lib.b200_gguf_arr_type.argtypes = [vp, i64]
    lib.b200_gguf_arr_n.restype = u64;           lib.b200_gguf_arr_n.argtypes = [vp, i64]
    lib.b200_gguf_val_data.restype = vp;         lib.b200_gguf_val_data.argtypes = [vp, i64]
    lib.b200_gguf_val_str.restype = C.c_char_p;  lib.b200_gguf_val_str.argtypes = [vp, i64, u64]
    lib.b200_gguf_val_number.restype = i32
    lib.b200_gguf_val_number.argtypes = [vp, i64, C.POINTER(C.c_double), C.POINTER(i64)]
    lib.b200_gguf_n_tensors.restype = i64;       lib.b200_gguf_n_tensors.argtypes = [vp]
    lib.b200_gguf_find_tensor.restype = i64;     lib.b200_gguf_find_tensor.argtypes = [vp, C.c_char_p]
    lib.b200_gguf_tensor.restype = i32;          lib.b200_gguf_tensor.argtypes = [vp, i64, C.POINTER(GgufTensorInfo)]
    lib.b200_gguf_type_row_bytes.restype = sz;   lib.b200_gguf_type_row_bytes.argtypes = [i32, i64]
    lib.b200_gguf_device_bytes.restype = sz;     lib.b200_gguf_device_bytes.argtypes = [vp]
    lib.b200_gguf_load.restype = i32;            lib.b200_gguf_load.argtypes = [vp, vp, sz, vp]
    lib.b200_gguf_load_tensors.restype = i32
    lib.b200_gguf_load_tensors.argtypes = [vp, C.POINTER(i64), C.POINTER(vp), i64, vp]
    return lib


class Gguf:
    """b200_gguf_* (include/ggml_b200.h): the reader half needs no GPU.  Mirrors gguf_init_from_file(no_alloc) + the
    gguf_get_* getters of the reference (include/gguf.h:78-128); a refused file raises B200Error with the reader's text."""
    _NP = {0: "<u1", 1: "<i1", 2: "<u2", 3: "<i2", 4: "<u4", 5: "<i4", 6: "<f4", 7: "<i1", 10: "<u8", 11: "<i8", 12: "<f8"}

    def __init__(self, path: str, lib: Optional[C.CDLL] = None):
        self.lib = lib or load_library()
        h = C.c_void_p()
        rc = self.lib.b200_gguf_open(os.fsencode(path), C.byref(h))
        if rc != 0:
            raise B200Error(f"gguf_open({path}): {self.lib.b200_gguf_last_error().decode(errors='replace')} (code {rc})")
        self.h = h

    def close(self):
        if self.h:
            self.lib.b200_gguf_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @property
    def version(self): return int(self.lib.b200_gguf_version(self.h))
    @property
    def alignment(self): return int(self.lib.b200_gguf_alignment(self.h))
    @property
    def data_offset(self): return int(self.lib.b200_gguf_data_offset(self.h))
    @property
    def data_bytes(self): return int(self.lib.b200_gguf_data_bytes(self.h))
    @property
    def device_bytes(self): return int(self.lib.b200_gguf_device_bytes(self.h))

    def keys(self):
        return [self.lib.b200_gguf_key(self.h, i).decode() for i in range(self.lib.b200_gguf_n_kv(self.h))]

    def value(self, key: str):
        """python value of a key: int / float / bool / str, or a numpy array / list of str for arrays"""
        import numpy as np
        i = self.lib.b200_gguf_find_key(self.h, key.encode())
        if i < 0:
            raise KeyError(key)
        t = self.lib.b200_gguf_kv_type(self.h, i)
        if t == 9:
            et, n = self.lib.b200_gguf_arr_type(self.h, i), int(self.lib.b200_gguf_arr_n(self.h, i))
            if et == 8:
                return [self.lib.b200_gguf_val_str(self.h, i, j).decode() for j in range(n)]
            dt = np.dtype(self._NP[et])
            raw = C.string_at(self.lib.b200_gguf_val_data(self.h, i), n * dt.itemsize) if n else b""
            return np.frombuffer(raw, dtype=dt).copy()
        if t == 8:
            return self.lib.b200_gguf_val_str(self.h, i, 0).decode()
        d, v = C.c_double(), C.c_int64()
        rc = self.lib.b200_gguf_val_number(self.h, i, C.byref(d), C.byref(v))
        if rc != 0:
            raise B200Error(f"gguf value {key}: code {rc}")
        return bool(v.value) if t == 7 else (float(d.value) if t in (6, 12) else int(v.value))

    def tensors(self):
        """[dict(name, type, n_dims, ne, file_offset, nbytes, device_offset)] in file order"""
        out = []
        ti = GgufTensorInfo()
        for i in range(self.lib.b200_gguf_n_tensors(self.h)):
            rc = self.lib.b200_gguf_tensor(self.h, i, C.byref(ti))
            if rc != 0:
                raise B200Error(f"gguf_tensor({i}): code {rc}")
            out.append(dict(name=ti.name.decode(), type=int(ti.type), n_dims=int(ti.n_dims), ne=[int(x) for x in ti.ne],
                            file_offset=int(ti.file_offset), nbytes=int(ti.nbytes), device_offset=int(ti.device_offset)))
        return out

    def find_tensor(self, name: str) -> int:
        return int(self.lib.b200_gguf_find_tensor(self.h, name.encode()))

    def load(self, kern: "Kernels"):
        """every tensor into one fresh device buffer (uint8); returns it (tensor i at tensors()[i]['device_offset'])"""
        torch = kern.torch
        buf = torch.empty(max(self.device_bytes, 256), dtype=torch.uint8, device=kern.device)
        kern._check(self.lib.b200_gguf_load(self.h, buf.data_ptr(), buf.numel(), kern._stream()), "gguf_load")
        return buf

    def load_tensors(self, kern: "Kernels", ids, dsts):
        n = len(ids)
        kern._check(self.lib.b200_gguf_load_tensors(self.h, (C.c_int64 * n)(*ids), (C.c_void_p * n)(*[int(d) for d in dsts]), n,
                                                    kern._stream()), "gguf_load_tensors")


def batch_plan(lib, nodes, count: Optional[int] = None):
    """[(launch group, waits-for list index or -1, route)] per node: how b200_mul_mat_batch would issue the list."""
    count = len(nodes) if count is None else count
    out = (C.c_int32 * (3 * max(count, 1)))()
    rc = lib.b200_mul_mat_batch_plan(nodes, count, out)
    if rc != 0:
        raise B200Error(f"mul_mat_batch_plan: code {rc}")
    return [(out[3 * i], out[3 * i + 1], out[3 * i + 2]) for i in range(count)]


def _a4i(v):
    return (C.c_int64 * 4)(*[int(x) for x in v])


def _a4s(v):
    return (C.c_size_t * 4)(*[int(x) for x in v])


class Kernels:
    """The kernel library bound to torch device buffers.  All methods enqueue on torch's current stream."""

    def __init__(self, path: Optional[str] = None, device: int = 0):
        import torch
        self.torch = torch
        self.lib = load_library(path)
        if not torch.cuda.is_available() or self.lib.b200_device_count() <= device:
            raise B200Error("no sm_100 CUDA device visible: the B200 kernels have no CPU fallback")
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(self.device)
        self._ws = None

    # ------------------------------------------------------------------ helpers
    def _check(self, rc: int, what: str):
        if rc != 0:
            raise B200Error(f"{what}: {self.lib.b200_error_string(rc).decode()} (code {rc})")

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def launch_count(self) -> int:
        return int(self.lib.b200_launch_count())

    def set_option(self, name: str, value: int) -> int:
        """b200_set_option: force one of the candidate kernels (tests / benches); returns the previous value."""
        return int(self.lib.b200_set_option(OPT[name], int(value)))

    def option(self, name: str, value: int):
        """context manager: `with kern.option("mmvq_generic", 1): ...`"""
        kern = self

        class _Ctx:
            def __enter__(self_inner):
                self_inner.prev = kern.set_option(name, value)

            def __exit__(self_inner, *exc):
                kern.set_option(name, self_inner.prev)
        return _Ctx()

    def tensor_peak(self, kind: str, iters: int = 4096, reps: int = 5) -> float:
        """ops/s (2 x MACs) of the tcgen05 micro-kernel: kind "i8" or "bf16" (include/ggml_b200.h: b200_tensor_peak)."""
        torch = self.torch
        a = torch.randint(0, 127, (128, 128), dtype=torch.uint8, device=self.device)
        b = torch.randint(0, 127, (256, 128), dtype=torch.uint8, device=self.device)
        out = C.c_double(0.0)
        self._check(self.lib.b200_tensor_peak(0 if kind == "i8" else 1, a.data_ptr(), b.data_ptr(), None, iters, reps,
                                              C.byref(out), self._stream()), "tensor_peak")
        return float(out.value)

    def tensor_peak_verify(self, kind: str, a, b):
        """one 128 x 256 x 128-byte product by the micro-kernel: a [128,128] / b [256,128] uint8 views -> [128,256] int32 / float32"""
        torch = self.torch
        out = torch.zeros((128, 256), dtype=torch.int32, device=self.device)
        self._check(self.lib.b200_tensor_peak(0 if kind == "i8" else 1, a.data_ptr(), b.data_ptr(), out.data_ptr(), 1, 1,
                                              None, self._stream()), "tensor_peak")
        return out if kind == "i8" else out.view(torch.float32)

    def device_error_flags(self, clear: bool = True) -> int:
        """sticky error bits set by kernels (bit 0: GET_ROWS index outside the table); synchronizes first."""
        self.torch.cuda.synchronize()
        return int(self.lib.b200_device_error_flags(1 if clear else 0))

    def to_device(self, arr):
        """numpy array (any dtype) -> device tensor of the same dtype/shape."""
        return self.torch.from_numpy(arr).to(self.device)

    def workspace(self, nbytes: int):
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = self.torch.empty(max(nbytes, 1 << 20), dtype=self.torch.uint8, device=self.device)
        return self._ws

    # ------------------------------------------------------------------ dequantize_row_*
    def dequantize_row(self, t: str, blocks, n_elems: int, out_dtype: str = "f32"):
        torch = self.torch
        dt = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}[out_dtype]
        out = torch.empty(n_elems, dtype=dt, device=self.device)
        self._check(self.lib.b200_dequantize_row(TYPE_ID[t], blocks.data_ptr(), out.data_ptr(), n_elems,
                                                 TYPE_ID[out_dtype], self._stream()), "dequantize_row")
        return out

    def dequantize_4d(self, t: str, blocks, ne_src, nb_src, out, ne_dst, nb_dst):
        out_type = TYPE_ID["f32"] if out.dtype == self.torch.float32 else TYPE_ID["f16"]
        self._check(self.lib.b200_dequantize_4d(TYPE_ID[t], blocks.data_ptr(), _a4i(ne_src), _a4s(nb_src), out.data_ptr(),
                                                out_type, _a4i(ne_dst), _a4s(nb_dst), self._stream()), "dequantize_4d")
        return out

    # ------------------------------------------------------------------ quantize_row_q8_1
    def quantize_row_q8_1(self, x):
        """x: f32 [n, k] (row stride may exceed k) -> uint8 [n, k/32*36] block_q8_1."""
        torch = self.torch
        assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
        n, k = x.shape
        out = torch.empty((n, k // 32 * 36), dtype=torch.uint8, device=self.device)
        self._check(self.lib.b200_quantize_row_q8_1(x.data_ptr(), x.stride(0) * 4, k, n, out.data_ptr(), self._stream()),
                    "quantize_row_q8_1")
        return out

    def quantize_rows(self, t: str, x):
        """x: f32 [n, k] -> uint8 [n, k/32 * block bytes] of Q8_0 / Q4_0 blocks (quantize_row_*_ref values)."""
        torch = self.torch
        n, k = x.shape
        out = torch.empty((n, row_bytes(t, k)), dtype=torch.uint8, device=self.device)
        self._check(self.lib.b200_quantize_rows(TYPE_ID[t], x.data_ptr(), x.stride(0) * 4, k, n, out.data_ptr(), out.stride(0), self._stream()),
                    "quantize_rows")
        return out

    def mul_mat_id(self, t: str, experts, m: int, k: int, n_as: int, x, ids):
        """MUL_MAT_ID: experts = n_as stacked matrices of m rows of k (block bytes, contiguous), x f32 [n_tok][ne11][k], ids int32
        [n_tok][n_used] -> f32 [n_tok][n_used][m]; slot e of token t multiplies expert ids[t, e] with x[t, e % ne11]."""
        torch = self.torch
        n_tok, ne11, _ = x.shape
        n_used = ids.shape[1]
        rb = row_bytes(t, k)
        ne_a, nb_a = _a4i([k, m, n_as, 1]), _a4s([BLOCK_BYTES[t], rb, rb * m, rb * m * n_as])
        ne_b, nb_b = _a4i([k, ne11, n_tok, 1]), _a4s([4, 4 * k, 4 * k * ne11, 4 * k * ne11 * n_tok])
        ne_i, nb_i = (C.c_int64 * 2)(n_used, n_tok), (C.c_size_t * 2)(4, 4 * n_used)
        ws = torch.empty(int(self.lib.b200_mul_mat_id_workspace_size(TYPE_ID[t], ne_a, ne_b)) + 256, dtype=torch.uint8, device=self.device)
        out = torch.empty((n_tok, n_used, m), dtype=torch.float32, device=self.device)
        self._check(self.lib.b200_mul_mat_id(TYPE_ID[t], experts.data_ptr(), ne_a, nb_a, x.data_ptr(), ne_b, nb_b, ids.data_ptr(), ne_i, nb_i,
                                             out.data_ptr(), ws.data_ptr(), ws.numel(), self._stream()), "mul_mat_id")
        return out

    def quantize_rows_4d(self, t: str, x, ne, nb_x, nb_y, out_bytes: int):
        """strided form: x f32 device tensor (any view), ne / nb_x / nb_y in ggml order (dim 0 first, strides in bytes)."""
        out = self.torch.zeros(out_bytes, dtype=self.torch.uint8, device=self.device)
        self._check(self.lib.b200_quantize_rows_4d(TYPE_ID[t], x.data_ptr(), _a4i(ne), _a4s(nb_x), out.data_ptr(), _a4s(nb_y), self._stream()),
                    "quantize_rows_4d")
        return out

    def quantize_acts_record(self, x):
        """x: f32 [k] -> the decode list kernel's q8 record (uint8, layout: include/ggml_b200.h)."""
        torch = self.torch
        k = x.numel()
        out = torch.zeros(int(self.lib.b200_acts_record_bytes(k)), dtype=torch.uint8, device=self.device)
        self._check(self.lib.b200_quantize_acts_record(x.data_ptr(), k, out.data_ptr(), self._stream()), "quantize_acts_record")
        return out

    def quantize_acts_columns(self, x):
        """x: f32 [n, k], n <= 8 -> the MMA list kernel's activation columns (uint8, layout: include/ggml_b200.h)."""
        torch = self.torch
        n, k = x.shape
        out = torch.zeros(int(self.lib.b200_acts_columns_bytes(k, n)), dtype=torch.uint8, device=self.device)
        self._check(self.lib.b200_quantize_acts_columns(x.data_ptr(), x.stride(0) * 4, k, n, out.data_ptr(), self._stream()),
                    "quantize_acts_columns")
        return out

    # ------------------------------------------------------------------ GET_ROWS
    def get_rows(self, t: str, table, k: int, rows, out_dtype: str = "f32", table_row_stride: Optional[int] = None):
        torch = self.torch
        assert rows.dtype == torch.int32
        rb = table_row_stride or row_bytes(t, k)
        n_table = table.numel() // rb
        out = torch.empty((rows.numel(), k), dtype=torch.float32 if out_dtype == "f32" else torch.float16, device=self.device)
        self._check(self.lib.b200_get_rows(TYPE_ID[t], table.data_ptr(), rb, k, n_table, rows.data_ptr(), 4, rows.numel(),
                                           out.data_ptr(), TYPE_ID[out_dtype], out.stride(0) * out.element_size(),
                                           self._stream()), "get_rows")
        return out

    # ------------------------------------------------------------------ MUL_MAT
    def quantize_acts(self, x, scratch=None):
        torch = self.torch
        n, k = x.shape
        nbytes = int(self.lib.b200_acts_scratch_bytes(k, n))
        if scratch is None:
            scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=self.device)
        self._check(self.lib.b200_quantize_acts(x.data_ptr(), x.stride(0) * 4, k, n, scratch.data_ptr(), self._stream()),
                    "quantize_acts")
        return scratch

    def mul_mat_vec_q(self, t: str, w, m: int, k: int, scratch, n: int, out=None, w_row_stride: Optional[int] = None):
        torch = self.torch
        if out is None:
            out = torch.empty((n, m), dtype=torch.float32, device=self.device)
        self._check(self.lib.b200_mul_mat_vec_q(TYPE_ID[t], w.data_ptr(), w_row_stride or row_bytes(t, k), m, k,
                                                scratch.data_ptr(), n, out.data_ptr(), out.stride(0) * 4, self._stream()),
                    "mul_mat_vec_q")
        return out

    def mul_mat_vec_static(self, t: str, w, m: int, k: int, x, out, w_row_stride: Optional[int] = None) -> int:
        """b200_mul_mat_vec_static: quantize + matvec on the library's static device scratch; allocates nothing (the
        caller passes `out` [n, m] f32).  Returns the C return code (0, or B200_ERR_TOO_LARGE = 7 ..) unchecked."""
        n = x.shape[0]
        return int(self.lib.b200_mul_mat_vec_static(TYPE_ID[t], w.data_ptr(), w_row_stride or row_bytes(t, k), m, k,
                                                    x.data_ptr(), x.stride(0) * 4, n, out.data_ptr(), out.stride(0) * 4,
                                                    self._stream()))

    def mul_mat_q(self, t: str, w, m: int, k: int, x, out=None, w_row_stride: Optional[int] = None):
        torch = self.torch
        n = x.shape[0]
        if out is None:
            out = torch.empty((n, m), dtype=torch.float32, device=self.device)
        nbytes = int(self.lib.b200_mul_mat_q_workspace_size(m, k, n))
        ws = self.workspace(nbytes)
        self._check(self.lib.b200_mul_mat_q(TYPE_ID[t], w.data_ptr(), w_row_stride or row_bytes(t, k), m, k,
                                            x.data_ptr(), x.stride(0) * 4, n, out.data_ptr(), out.stride(0) * 4,
                                            ws.data_ptr(), ws.numel(), self._stream()), "mul_mat_q")
        return out

    def mul_mat(self, t: str, w, m: int, k: int, x, out=None, const_weights: bool = False, w_row_stride: Optional[int] = None):
        """ggml_mul_mat(W[k,m] quantized, X[k,n] f32) -> [m,n] stored as an (n, m) row-major tensor."""
        n = x.shape[0]
        ne_a = [k, m, 1, 1]
        rb = w_row_stride or row_bytes(t, k)
        nb_a = [BLOCK_BYTES[t], rb, rb * m, rb * m]
        ne_b = [k, n, 1, 1]
        sb = x.stride(0) * 4
        nb_b = [4, sb, sb * n, sb * n]
        return self.mul_mat_4d(t, w, ne_a, nb_a, x, ne_b, nb_b, out, flags=1 if const_weights else 0)

    def mul_mat_4d(self, t: str, a, ne_a, nb_a, b, ne_b, nb_b, out=None, flags: int = 0):
        torch = self.torch
        if out is None:
            out = torch.empty((ne_b[3], ne_b[2], ne_b[1], ne_a[1]), dtype=torch.float32, device=self.device)
        nbytes = int(self.lib.b200_mul_mat_workspace_size(TYPE_ID[t], _a4i(ne_a), _a4i(ne_b)))
        ws = self.workspace(nbytes)
        self._check(self.lib.b200_mul_mat_ex(TYPE_ID[t], a.data_ptr(), _a4i(ne_a), _a4s(nb_a), b.data_ptr(), _a4i(ne_b),
                                             _a4s(nb_b), out.data_ptr(), ws.data_ptr(), ws.numel(), flags, self._stream()),
                    "mul_mat")
        return out

    def mul_mat_batch_peers(self, nodes, peers: PeerGroup, count: Optional[int] = None, ws=None):
        """row split: the list on this rank's row blocks with the gather fused into the launch (every rank calls it)."""
        count = len(nodes) if count is None else count
        if ws is None:
            ws = self.workspace(int(self.lib.b200_mul_mat_batch_workspace_size(nodes, count)))
        self._check(self.lib.b200_mul_mat_batch_peers(nodes, count, C.byref(peers), ws.data_ptr(), ws.numel(), self._stream()),
                    "mul_mat_batch_peers")

    def mul_mat_batch(self, nodes, count: Optional[int] = None, ws=None):
        """nodes: a ctypes array of MulMatNode (see mul_mat_node); one host call for the whole list."""
        count = len(nodes) if count is None else count
        if ws is None:
            ws = self.workspace(int(self.lib.b200_mul_mat_batch_workspace_size(nodes, count)))
        self._check(self.lib.b200_mul_mat_batch(nodes, count, ws.data_ptr(), ws.numel(), self._stream()), "mul_mat_batch")
