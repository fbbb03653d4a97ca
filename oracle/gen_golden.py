#!/usr/bin/env python3
"""Generate tests/golden/golden_v1.npz FROM THE UNMODIFIED REFERENCE (oracle/_ref, built by
oracle/Makefile from /root/reference).  TEST INFRASTRUCTURE ONLY.

The reference ships no fixed-value vectors for this path (test-backend-ops draws from
std::random_device and compares live against ggml-cpu), so these are "outputs of the reference
itself run here", produced by calling the reference's own exported functions:

  * ggml_quantize_chunk            (src/ggml.c, include/ggml.h:2409)   — random-quantized weights,
                                   same recipe as tests/test-backend-ops.cpp:83-119
  * dequantize_row_{q4_0,q8_0,q4_K,q5_K,q6_K} (src/ggml-quants.c:307,401,1352,1554,1762)
  * quantize_row_q8_1_ref          (src/ggml-quants.c:225-258)
  * ggml_fp16_to_fp32_row / ggml_fp32_to_fp16_row (src/ggml.c:455-490)
  * ggml_mul_mat on the ggml-cpu backend (src/ggml-cpu/ggml-cpu.c:1202-1394)
  * GET_ROWS == dequantize of the selected rows (src/ggml-cpu/ops.cpp:5121-5165)

Run here (needs oracle/_ref):   python oracle/gen_golden.py
The .npz is committed; the GPU box never needs /root/reference.
"""
import hashlib
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("ggml_host", os.path.join(ROOT, "tests", "ggml_host.py"))
gh = importlib.util.module_from_spec(spec)
spec.loader.exec_module(gh)

K = 512           # elements per row in the block fixtures (2 super-blocks / 16 small blocks)
F16_FIELDS = {    # byte offsets of the fp16 fields inside a block
    "q4_0": (0,), "q8_0": (0,), "q4_K": (0, 2), "q5_K": (0, 2), "q6_K": (208,),
}
EDGE_F16 = [0x0000, 0x8000, 0x0001, 0x03FF, 0x0400, 0x3C00, 0xBC00, 0x7BFF, 0xFBFF, 0x1234, 0x3555, 0x6789]


def adversarial_blocks(tname, rng):
    """Raw random bytes as blocks (every bit pattern of the integer fields is legal), with the fp16
    scale fields forced through the edge values (zero, -0, denormals, max finite, ...)."""
    bs = gh.BLOCK_BYTES[tname]
    nblk_row = K // gh.BLOCK_ELEMS[tname]
    nrows = 6
    raw = rng.integers(0, 256, size=(nrows, nblk_row, bs), dtype=np.uint8)
    i = 0
    for r in range(nrows):
        for b in range(nblk_row):
            for off in F16_FIELDS[tname]:
                v = EDGE_F16[i % len(EDGE_F16)]
                if r >= 4:   # last rows: random finite halves
                    v = int(rng.integers(0, 0x7C00)) | (int(rng.integers(0, 2)) << 15)
                raw[r, b, off] = v & 0xFF
                raw[r, b, off + 1] = v >> 8
                i += 1
    if tname in ("q4_K", "q5_K"):
        raw[0, 0, 4:16] = 0xFF      # all 6-bit scales and mins = 63
        raw[0, 1, 4:16] = 0x00
    if tname == "q6_K":
        raw[0, 0, 192:208] = 0x80   # int8 scales = -128
        raw[0, 1, 192:208] = 0x7F
    return raw.reshape(nrows, nblk_row * bs)


def main():
    host = gh.GgmlHost(os.path.join(ROOT, "oracle", "_ref", "bin"))
    rng = np.random.default_rng(0xB200)
    out = {}

    # ---- weights: random-quantized (harness recipe) + adversarial raw blocks, and their dequantization
    for t in gh.QUANT_TYPES:
        x = rng.uniform(-1.0, 1.0, size=(6, K)).astype(np.float32)
        x[2] *= 1e-3          # tiny magnitudes
        x[3, ::7] = 0.0
        q_plain = host.quantize(t, x[:3], None)
        q_imat = host.quantize(t, x[3:], np.ones(K, dtype=np.float32))
        adv = adversarial_blocks(t, rng)
        blocks = np.concatenate([q_plain, q_imat, adv], axis=0)
        deq = np.stack([host.dequantize(t, blocks[r], K) for r in range(blocks.shape[0])])
        assert np.isfinite(deq).all(), t
        out[f"{t}_src_f32"] = x
        out[f"{t}_blocks"] = blocks
        out[f"{t}_deq_f32"] = deq
        out[f"{t}_deq_f16"] = host.fp32_to_fp16(deq.reshape(-1)).reshape(deq.shape)
        # GET_ROWS: a gather with repeats and reversed order
        rows = np.array([5, 0, 0, 11, 3, 7, 11], dtype=np.int32)
        out[f"{t}_getrows_idx"] = rows
        out[f"{t}_getrows_f32"] = deq[rows]

    # ---- q8_1 activation quantization
    xs = []
    xs.append(rng.uniform(-1, 1, size=1024).astype(np.float32))
    xs.append((rng.standard_normal(1024) * 10).astype(np.float32))
    xs.append(np.zeros(64, dtype=np.float32))
    o = rng.uniform(-1e-3, 1e-3, size=256).astype(np.float32); o[17] = 300.0; o[100] = -4e4; xs.append(o)
    # exact ties: x = (j + 0.5) * d with amax = 127 d  -> q = round-half-away-from-zero
    d = np.float32(0.25)
    tie = np.array([(j + 0.5) * d * (1 if j % 2 else -1) for j in range(31)] + [127 * d], dtype=np.float32)
    xs.append(np.tile(tie, 4))
    xs.append((rng.uniform(-1, 1, size=256) * 1e-41).astype(np.float32))   # f32 denormals
    xs.append((rng.uniform(-1, 1, size=256) * 3e38).astype(np.float32))    # near f32 max
    xs.append((rng.uniform(-1, 1, size=256) * 7e4).astype(np.float32))     # d overflows... not: amax/127 < 65504
    xs.append((rng.uniform(-1, 1, size=256) * 1e7).astype(np.float32))     # d > fp16 max -> inf in the block
    x81 = np.concatenate(xs)
    out["q8_1_x"] = x81
    out["q8_1_blocks"] = host.quantize_q8_1(x81)

    # ---- fp16 conversions: all 65536 halves -> f32 (digest), and f32 -> f16 on edges + random bits
    allh = np.arange(65536, dtype=np.uint16)
    f = host.fp16_to_fp32(allh)
    out["fp16_all_to_f32_sha256"] = np.frombuffer(hashlib.sha256(f.tobytes()).digest(), dtype=np.uint8)
    edges = np.array([0.0, -0.0, 5.9604645e-08, 2.9802322e-08, 2.9802326e-08, 8.940697e-08, 6.0975552e-05,
                      6.1035156e-05, 6.1005354e-05, 1.0, 1.00048828125, 1.000488281, 1.00146484375, 65504.0, 65519.99,
                      65520.0, 65536.0, 1e10, -1e10, np.inf, -np.inf, 1e-10, 0.333333343], dtype=np.float32)
    rb = rng.integers(0, 2**32, size=1 << 16, dtype=np.uint64).astype(np.uint32).view(np.float32)
    rb = rb[~np.isnan(rb)]
    sm = (rng.uniform(-1, 1, size=1 << 14) * np.exp(rng.uniform(-20, 12, size=1 << 14))).astype(np.float32)
    f32in = np.concatenate([edges, rb, sm])
    out["f32_to_f16_in"] = f32in
    out["f32_to_f16_out"] = host.fp32_to_fp16(f32in)

    # ---- MUL_MAT through the reference CPU backend
    cpu = host.cpu_backend(4)
    m = 32
    X = rng.uniform(-1, 1, size=(16, K)).astype(np.float32)
    out["mm_x"] = X
    for t in gh.QUANT_TYPES:
        w = rng.uniform(-1, 1, size=(m, K)).astype(np.float32)
        wq = host.quantize(t, w, None)
        out[f"mm_{t}_w"] = wq
        for n in (1, 3, 8, 16):
            out[f"mm_{t}_y_n{n}"] = host.mul_mat(cpu, t, wq, m, K, X[:n])

    dst = os.path.join(ROOT, "tests", "golden")
    os.makedirs(dst, exist_ok=True)
    path = os.path.join(dst, "golden_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")


def main_v2():
    """tests/golden/golden_v2.npz: quantize_row_q8_K_ref (src/ggml-quants.c:2555-2592) — the activation format of ggml-cpu's
    K-quant MUL_MAT, which the 2..8-column decode kernel uses too.  Separate file: golden_v1.npz is left as committed."""
    host = gh.GgmlHost(os.path.join(ROOT, "oracle", "_ref", "bin"))
    rng = np.random.default_rng(0xB201)
    xs = [rng.uniform(-1, 1, size=1024).astype(np.float32), (rng.standard_normal(1024) * 10).astype(np.float32)]
    z = rng.uniform(-1, 1, size=768).astype(np.float32); z[256:512] = 0.0; xs.append(z)          # an all-zero super-block
    o = rng.uniform(-1e-3, 1e-3, size=256).astype(np.float32); o[17] = 300.0; xs.append(o)         # outlier (positive max -> -127)
    o = rng.uniform(-1e-3, 1e-3, size=256).astype(np.float32); o[100] = -4e4; xs.append(o)         # negative max -> +127 ... min(127, v)
    t = rng.uniform(-0.5, 0.5, size=256).astype(np.float32); t[9] = 1.0; t[200] = -1.0; xs.append(t)   # |max| tie: the FIRST one counts
    t = rng.uniform(-0.5, 0.5, size=256).astype(np.float32); t[9] = -1.0; t[200] = 1.0; xs.append(t)
    tie = np.array([(j + 0.5) / 127.0 * (1 if j % 2 else -1) for j in range(255)] + [-1.0], dtype=np.float32); xs.append(tie)  # x.5 ties
    xs.append((rng.uniform(-1, 1, size=256) * 1e-30).astype(np.float32))
    xs.append((rng.uniform(-1, 1, size=256) * 1e30).astype(np.float32))
    x = np.concatenate(xs)
    q, d, bs = host.quantize_q8_K(x)
    zero = d == 0
    bs = bs.reshape(-1, 16).copy(); bs[zero] = 0        # the reference leaves bsums of an all-zero block unwritten
    out = {"q8_K_x": x, "q8_K_q": q, "q8_K_d": d, "q8_K_bsums": bs.reshape(-1)}
    # the formats only the general kernels serve: random-quantized rows + raw random blocks with edge fp16 fields, dequantized
    for t in gh.EXTRA_TYPES:
        F16_FIELDS[t] = {"q4_1": (0, 2), "q5_0": (0,), "q5_1": (0, 2), "iq4_nl": (0,)}[t]
        xw = rng.uniform(-1.0, 1.0, size=(4, K)).astype(np.float32)
        xw[2] *= 1e-3
        blocks = np.concatenate([host.quantize(t, xw, None), adversarial_blocks(t, rng)], axis=0)
        deq = np.stack([host.dequantize(t, blocks[r], K) for r in range(blocks.shape[0])])
        out[f"{t}_blocks"] = blocks
        out[f"{t}_deq_f32"] = deq
        out[f"{t}_deq_f16"] = host.fp32_to_fp16(deq.reshape(-1)).reshape(deq.shape)
    path = os.path.join(ROOT, "tests", "golden", "golden_v2.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    sys.exit(main_v2() if "v2" in sys.argv[1:] else main())
