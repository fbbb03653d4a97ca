"""ctypes binding of oracle/libb200_oracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this.  The product (ggml-gfx906_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

TYPE_ID = {"f32": 0, "f16": 1, "q4_0": 2, "q8_0": 8, "q8_1": 9, "q4_K": 12, "q5_K": 13, "q6_K": 14,
           "q4_1": 3, "q5_0": 6, "q5_1": 7, "iq4_nl": 20}
BLOCK_ELEMS = {"q4_0": 32, "q8_0": 32, "q8_1": 32, "q4_K": 256, "q5_K": 256, "q6_K": 256, "q4_1": 32, "q5_0": 32, "q5_1": 32, "iq4_nl": 32}
BLOCK_BYTES = {"q4_0": 18, "q8_0": 34, "q8_1": 36, "q4_K": 144, "q5_K": 176, "q6_K": 210, "q4_1": 20, "q5_0": 22, "q5_1": 24, "iq4_nl": 18}
QUANT_TYPES = ("q4_0", "q8_0", "q4_K", "q5_K", "q6_K")           # the hot path: every kernel
EXTRA_TYPES = ("q4_1", "q5_0", "q5_1", "iq4_nl")                 # dequantize / GET_ROWS / CPY and the general mat-vec only


def build():
    so = os.path.join(ORACLE_DIR, "libb200_oracle.so")
    src = os.path.join(ORACLE_DIR, "b200_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "oracle"], stdout=subprocess.DEVNULL)
    return so


class Oracle:
    def __init__(self):
        self.lib = C.CDLL(build())
        L, vp, i64, sz = self.lib, C.c_void_p, C.c_int64, C.c_size_t
        L.b200o_dequantize_row.argtypes = [C.c_int, vp, vp, i64]
        L.b200o_dequantize_row_f16.argtypes = [C.c_int, vp, vp, i64]
        L.b200o_quantize_row_q8_1.argtypes = [vp, vp, i64]
        L.b200o_quantize_row_q8_0.argtypes = [vp, vp, i64]
        L.b200o_quantize_row_q8_K.argtypes = [vp, vp, vp, vp, i64]
        L.b200o_quantize_row_q4_0.argtypes = [vp, vp, i64]
        L.b200o_get_rows.argtypes = [C.c_int, vp, sz, i64, vp, i64, vp]
        L.b200o_mul_mat.argtypes = [C.c_int, vp, sz, i64, i64, vp, i64, vp, C.c_int]
        L.b200o_nmse.argtypes = [vp, vp, sz]
        L.b200o_nmse.restype = C.c_double
        L.b200o_fp16_to_fp32.argtypes = [C.c_uint16]
        L.b200o_fp16_to_fp32.restype = C.c_float
        L.b200o_fp32_to_fp16.argtypes = [C.c_float]
        L.b200o_fp32_to_fp16.restype = C.c_uint16

    @staticmethod
    def row_bytes(t, k):
        assert k % BLOCK_ELEMS[t] == 0
        return k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]

    def dequantize(self, t, blocks, k_total):
        blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
        out = np.empty(k_total, dtype=np.float32)
        self.lib.b200o_dequantize_row(TYPE_ID[t], blocks.ctypes.data, out.ctypes.data, k_total)
        return out

    def dequantize_f16(self, t, blocks, k_total):
        blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
        out = np.empty(k_total, dtype=np.uint16)
        self.lib.b200o_dequantize_row_f16(TYPE_ID[t], blocks.ctypes.data, out.ctypes.data, k_total)
        return out

    def quantize_q8_1(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
        out = np.empty(x.size // 32 * 36, dtype=np.uint8)
        self.lib.b200o_quantize_row_q8_1(x.ctypes.data, out.ctypes.data, x.size)
        return out

    def quantize_q8_K(self, x):
        """quantize_row_q8_K_ref -> (q int8 [k], d f32 [k/256], bsums int16 [k/16])"""
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
        q = np.empty(x.size, dtype=np.int8)
        d = np.empty(x.size // 256, dtype=np.float32)
        bs = np.empty(x.size // 16, dtype=np.int16)
        self.lib.b200o_quantize_row_q8_K(x.ctypes.data, q.ctypes.data, d.ctypes.data, bs.ctypes.data, x.size)
        return q, d, bs

    def quantize(self, t, x):
        """q8_0 / q4_0 only (the K-quant quantizers are not on the hot path; see random_blocks)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        nrows, k = x.shape
        out = np.empty((nrows, self.row_bytes(t, k)), dtype=np.uint8)
        fn = {"q8_0": self.lib.b200o_quantize_row_q8_0, "q4_0": self.lib.b200o_quantize_row_q4_0}[t]
        fn(x.ctypes.data, out.ctypes.data, nrows * k)
        return out

    def get_rows(self, t, table, k, rows):
        table = np.ascontiguousarray(table, dtype=np.uint8)
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        out = np.empty((rows.size, k), dtype=np.float32)
        self.lib.b200o_get_rows(TYPE_ID[t], table.ctypes.data, self.row_bytes(t, k), k,
                                rows.ctypes.data, rows.size, out.ctypes.data)
        return out

    def mul_mat(self, t, w, m, k, x, mode=1, row_stride=None):
        w = np.ascontiguousarray(w, dtype=np.uint8)
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, k)
        n = x.shape[0]
        out = np.empty((n, m), dtype=np.float32)
        self.lib.b200o_mul_mat(TYPE_ID[t], w.ctypes.data, row_stride or self.row_bytes(t, k), m, k,
                               x.ctypes.data, n, out.ctypes.data, mode)
        return out

    def nmse(self, a, b):
        a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
        b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1)
        assert a.size == b.size
        return float(self.lib.b200o_nmse(a.ctypes.data, b.ctypes.data, a.size))


# ---------------------------------------------------------------------------------------------
# Seeded synthetic weights that need neither the reference nor a K-quant quantizer: every bit
# pattern of the integer fields of a block is legal, so random bytes with sane fp16 scale fields are
# valid "random-quantized" weights.  Scales are drawn so that dequantized values are O(1).
F16_FIELDS = {"q4_0": (0,), "q8_0": (0,), "q4_K": (0, 2), "q5_K": (0, 2), "q6_K": (208,),
              "q4_1": (0, 2), "q5_0": (0,), "q5_1": (0, 2), "iq4_nl": (0,)}
_SCALE_RANGE = {"q4_0": (0.02, 0.15), "q8_0": (0.001, 0.01), "q4_K": (0.0005, 0.004),
                "q5_K": (0.0003, 0.002), "q6_K": (0.0001, 0.0006),
                "q4_1": (0.02, 0.1), "q5_0": (0.01, 0.06), "q5_1": (0.01, 0.05), "iq4_nl": (0.002, 0.008)}


def random_blocks(t, nrows, k, seed):
    rng = np.random.default_rng(seed)
    nb = k // BLOCK_ELEMS[t]
    bs = BLOCK_BYTES[t]
    raw = rng.integers(0, 256, size=(nrows, nb, bs), dtype=np.uint8)
    lo, hi = _SCALE_RANGE[t]
    for off in F16_FIELDS[t]:
        s = rng.uniform(lo, hi, size=(nrows, nb)).astype(np.float16)
        if off == 0 or off == 208:
            s = s * rng.choice(np.array([-1, 1], dtype=np.float16), size=s.shape)
        v = s.view(np.uint16)
        raw[:, :, off] = (v & 0xFF).astype(np.uint8)
        raw[:, :, off + 1] = (v >> 8).astype(np.uint8)
    return raw.reshape(nrows, nb * bs)


def scale_blocks(t, blocks, factor):
    """multiply every fp16 scale field of the blocks by `factor` (in place, RN to fp16): dequantized values scale with it"""
    nrows = blocks.shape[0]
    raw = blocks.reshape(nrows, -1, BLOCK_BYTES[t])
    for off in F16_FIELDS[t]:
        v = (raw[:, :, off].astype(np.uint16) | (raw[:, :, off + 1].astype(np.uint16) << 8)).view(np.float16)
        v = (v.astype(np.float32) * np.float32(factor)).astype(np.float16).view(np.uint16)
        raw[:, :, off] = (v & 0xFF).astype(np.uint8)
        raw[:, :, off + 1] = (v >> 8).astype(np.uint8)
    return blocks
