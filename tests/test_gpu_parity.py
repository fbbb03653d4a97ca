"""GPU parity tests: the CUDA path, called through the C ABI (include/ggml_b200.h), against the CPU
oracle (oracle/b200_oracle.c) on the same seeded inputs, and against the committed golden vectors
the unmodified reference produced (tests/golden/golden_v1.npz).

Bars (BASELINE.json north_star): dequantization, GET_ROWS and q8_1 activation quantization are
BIT-EXACT; MUL_MAT is within test-backend-ops' NMSE gate 5e-4 of ggml-cpu (golden) and — much
tighter — within NMSE 1e-9 of the oracle's "mode 1" (weights dequantized, activations rounded
through q8_1: the arithmetic the kernel performs, differing only in f32 summation order).
"""
import os

import numpy as np
import pytest

from oracle_lib import BLOCK_BYTES, BLOCK_ELEMS, QUANT_TYPES, random_blocks

pytestmark = pytest.mark.gpu

NMSE_REF = 5e-4      # tests/test-backend-ops.cpp:3106-3108 (MUL_MAT max_nmse_err)
NMSE_ORACLE = 1e-9   # same arithmetic, different f32 summation order


def u32(a):
    return np.ascontiguousarray(a).view(np.uint32)


def nmse64(want, got):
    """NMSE in float64 (the oracle's helper squares in f32: values near 1e+-30 over/underflow there)"""
    w, g = np.asarray(want, dtype=np.float64).reshape(-1), np.asarray(got, dtype=np.float64).reshape(-1)
    return float(((w - g) ** 2).sum() / (w ** 2).sum())


def dev_blocks(kern, blocks):
    return kern.to_device(np.ascontiguousarray(blocks).reshape(-1))


def adversarial_blocks(t, nrows, k, seed):
    """random blocks with edge-case scale fields: zero, subnormal, max-normal, negative."""
    raw = random_blocks(t, nrows, k, seed).reshape(nrows, k // BLOCK_ELEMS[t], BLOCK_BYTES[t]).copy()
    specials = np.array([0x0000, 0x8000, 0x0001, 0x03FF, 0x0400, 0x7BFF, 0xFBFF, 0x3C00], dtype=np.uint16)
    offs = {"q4_0": (0,), "q8_0": (0,), "q4_K": (0, 2), "q5_K": (0, 2), "q6_K": (208,)}[t]
    rng = np.random.default_rng(seed + 1)
    for off in offs:
        pick = specials[rng.integers(0, len(specials), size=raw.shape[:2])]
        mask = rng.random(raw.shape[:2]) < 0.5
        lo = np.where(mask, (pick & 0xFF).astype(np.uint8), raw[:, :, off])
        hi = np.where(mask, (pick >> 8).astype(np.uint8), raw[:, :, off + 1])
        raw[:, :, off], raw[:, :, off + 1] = lo, hi
    if t in ("q4_K", "q5_K"):
        raw[0, :, 4:16] = 0xFF     # every 6-bit scale and min = 63
    if t == "q6_K":
        raw[0, :, 192:208] = 0x80  # every scale = -128
    return raw.reshape(nrows, -1)


# ---------------------------------------------------------------------------- dequantize_row_*
@pytest.mark.parametrize("t", QUANT_TYPES)
def test_dequantize_golden(kern, golden, t):
    blocks = golden[f"{t}_blocks"]
    got32 = kern.dequantize_row(t, dev_blocks(kern, blocks), blocks.shape[0] * 512, "f32").cpu().numpy()
    assert np.array_equal(u32(got32), u32(golden[f"{t}_deq_f32"].reshape(-1)))
    got16 = kern.dequantize_row(t, dev_blocks(kern, blocks), blocks.shape[0] * 512, "f16").cpu().numpy()
    assert np.array_equal(got16.view(np.uint16), golden[f"{t}_deq_f16"].reshape(-1))


@pytest.mark.parametrize("t", QUANT_TYPES)
@pytest.mark.parametrize("maker", ["random", "adversarial"])
def test_dequantize_vs_oracle(kern, oracle, t, maker):
    nrows, k = 64, 4096
    blocks = (random_blocks if maker == "random" else adversarial_blocks)(t, nrows, k, seed=11)
    want = oracle.dequantize(t, blocks, nrows * k)
    got = kern.dequantize_row(t, dev_blocks(kern, blocks), nrows * k, "f32").cpu().numpy()
    assert np.array_equal(u32(got), u32(want))
    want16 = oracle.dequantize_f16(t, blocks, nrows * k)
    got16 = kern.dequantize_row(t, dev_blocks(kern, blocks), nrows * k, "f16").cpu().numpy().view(np.uint16)
    assert np.array_equal(got16, want16)


@pytest.mark.parametrize("t", QUANT_TYPES)
def test_dequantize_edge_sizes(kern, oracle, t):
    # one block, and zero elements
    k = BLOCK_ELEMS[t]
    blocks = random_blocks(t, 1, k, seed=3)
    got = kern.dequantize_row(t, dev_blocks(kern, blocks), k, "f32").cpu().numpy()
    assert np.array_equal(u32(got), u32(oracle.dequantize(t, blocks, k)))
    assert kern.dequantize_row(t, dev_blocks(kern, blocks), 0, "f32").numel() == 0
    with pytest.raises(Exception):
        kern.dequantize_row(t, dev_blocks(kern, blocks), k - 1, "f32")


@pytest.mark.parametrize("t", QUANT_TYPES)
def test_cpy_q_to_f32_permuted(kern, oracle, t):
    """CPY(quant -> f32) with a permuted destination, the test-backend-ops CPY case shape {256,2,3,4}."""
    import torch
    ne = [256, 2, 3, 4]
    nrows = ne[1] * ne[2] * ne[3]
    blocks = random_blocks(t, nrows, ne[0], seed=5)
    rb = blocks.shape[1]
    nb_src = [BLOCK_BYTES[t], rb, rb * ne[1], rb * ne[1] * ne[2]]
    want = oracle.dequantize(t, blocks, nrows * ne[0]).reshape(ne[3], ne[2], ne[1], ne[0])
    # destination: same logical shape, dims 1 and 2 swapped in memory
    out = torch.zeros((ne[3], ne[1], ne[2], ne[0]), dtype=torch.float32, device=kern.device)
    nb_dst = [4, ne[0] * ne[2] * 4, ne[0] * 4, ne[0] * ne[1] * ne[2] * 4]
    kern.dequantize_4d(t, dev_blocks(kern, blocks), ne, nb_src, out, ne, nb_dst)
    got = out.cpu().numpy().transpose(0, 2, 1, 3)
    assert np.array_equal(u32(got), u32(want))


# ---------------------------------------------------------------------------- GET_ROWS
@pytest.mark.parametrize("t", QUANT_TYPES)
def test_get_rows_golden(kern, golden, t):
    import torch
    blocks = golden[f"{t}_blocks"]
    idx = torch.from_numpy(golden[f"{t}_getrows_idx"]).to(kern.device)
    got = kern.get_rows(t, dev_blocks(kern, blocks), 512, idx).cpu().numpy()
    assert np.array_equal(u32(got), u32(golden[f"{t}_getrows_f32"]))


@pytest.mark.parametrize("t", QUANT_TYPES)
def test_get_rows_embedding_table(kern, oracle, t):
    import torch
    n_vocab, k = 2000, 4096
    blocks = random_blocks(t, n_vocab, k, seed=21)
    rng = np.random.default_rng(22)
    rows = rng.integers(0, n_vocab, size=37).astype(np.int32)
    rows[:3] = [0, n_vocab - 1, 0]      # first, last, repeated
    got = kern.get_rows(t, dev_blocks(kern, blocks), k, torch.from_numpy(rows).to(kern.device)).cpu().numpy()
    assert np.array_equal(u32(got), u32(oracle.get_rows(t, blocks, k, rows)))


# ---------------------------------------------------------------------------- quantize_row_q8_1
def _act_cases():
    rng = np.random.default_rng(31)
    x = rng.uniform(-1, 1, size=(4, 4096)).astype(np.float32)
    xn = rng.standard_normal(size=(4, 4096)).astype(np.float32) * 3
    z = np.zeros((1, 4096), dtype=np.float32)
    out = rng.standard_normal(size=(2, 4096)).astype(np.float32)
    out[:, ::97] *= 1000.0
    # exact ties: values k + 0.5 in units of d (d = amax/127 with amax = 127 -> d = 1)
    ties = np.tile(np.concatenate([[127.0], np.arange(-15, 16) + 0.5]).astype(np.float32), (1, 128))
    tiny = (rng.standard_normal(size=(1, 4096)) * 1e-30).astype(np.float32)
    return {"uniform": x, "normal": xn, "zeros": z, "outliers": out, "ties": ties, "tiny": tiny}


@pytest.mark.parametrize("case", list(_act_cases().keys()))
def test_quantize_q8_1_vs_oracle(kern, oracle, case):
    x = _act_cases()[case]
    got = kern.quantize_row_q8_1(kern.to_device(x)).cpu().numpy()
    want = np.stack([oracle.quantize_q8_1(r) for r in x])
    assert np.array_equal(got, want)


def test_quantize_q8_1_golden(kern, golden):
    x = golden["q8_1_x"].reshape(1, -1)
    got = kern.quantize_row_q8_1(kern.to_device(x)).cpu().numpy().reshape(-1)
    assert np.array_equal(got, golden["q8_1_blocks"])


def test_quantize_q8_1_strided_rows(kern, oracle):
    rng = np.random.default_rng(33)
    big = rng.standard_normal(size=(5, 1024)).astype(np.float32)
    view = kern.to_device(big)[:, :512]          # row stride 1024 floats, k = 512
    got = kern.quantize_row_q8_1(view).cpu().numpy()
    want = np.stack([oracle.quantize_q8_1(r[:512]) for r in big])
    assert np.array_equal(got, want)


def _q8_1_reference(oracle, x):
    """(int8 values [k], fp16-rounded d as f32 [k/32], integer sums [k/32]) of quantize_row_q8_1_ref"""
    blocks = oracle.quantize_q8_1(x).reshape(-1, 36)
    d = blocks[:, 0:2].copy().view(np.float16).astype(np.float32).reshape(-1)
    q = blocks[:, 4:36].copy().view(np.int8).reshape(-1)
    return q, d, q.reshape(-1, 32).astype(np.int32).sum(axis=1)


@pytest.mark.parametrize("case", list(_act_cases().keys()))
def test_quantize_list_record_vs_oracle(kern, oracle, case):
    """k_quantize_batch, the quantizer on the headline path (decode list kernel): its record, un-swizzled on the host,
    is quantize_row_q8_1_ref bit for bit (ties, tiny values and outliers included)."""
    for x in _act_cases()[case][:2]:
        k = x.size
        rec = kern.quantize_acts_record(kern.to_device(x)).cpu().numpy()
        e = np.arange(k)
        idx, pc = e >> 7, (e >> 4) & 7
        pos = idx * 128 + (((pc + idx) & 7) << 4) + (e & 15)
        got_q = rec[:k].view(np.int8)[pos]
        got_d = rec[k:k + k // 32 * 4].view(np.float32)
        q, d, _ = _q8_1_reference(oracle, x)
        assert np.array_equal(got_q, q)
        assert np.array_equal(u32(got_d), u32(d))


@pytest.mark.parametrize("case", list(_act_cases().keys()))
@pytest.mark.parametrize("n", [2, 5, 8])
def test_quantize_list_columns_vs_oracle(kern, oracle, case, n):
    """k_quantize_cols (decode at batch 2..8): int8 [k/32][8][32], then per super-block the f32 scale of each column and the
    sums of 32 / of 16 quants as (sum & 63, sum >> 6) byte pairs -- bit for bit quantize_row_q8_K_ref
    (src/ggml-quants.c:2555-2592) per column; padding columns carry zero scales and sums."""
    base = _act_cases()[case]
    k = base.shape[1] // 256 * 256
    x = np.ascontiguousarray(np.resize(base[:, :k], (n, k)) * np.linspace(1.0, 2.0, n, dtype=np.float32)[:, None])
    if case == "ties":       # equal magnitudes of opposite sign inside a super-block: the first one sets the sign of the scale
        x[0, 5], x[0, 200] = -np.abs(x[0]).max() * 2, np.abs(x[0]).max() * 2
        x[1, 7], x[1, 100] = np.abs(x[1]).max() * 2, -np.abs(x[1]).max() * 2
    buf = kern.quantize_acts_columns(kern.to_device(x)).cpu().numpy()
    nsb = k // 256
    yq = buf[:8 * k].view(np.int8).reshape(k // 32, 8, 32)
    off = (8 * k + 127) & ~127
    d = buf[off:off + nsb * 32].view(np.float32).reshape(nsb, 8)
    s32 = buf[off + nsb * 32: off + nsb * 160].reshape(nsb, 8, 2, 2, 4)        # [sb][col][t][lo / hi][4 sub-blocks]
    s16 = buf[off + nsb * 160: off + nsb * 416].reshape(nsb, 8, 4, 2, 4)
    for c in range(n):
        q, dq, bs = oracle.quantize_q8_K(x[c])
        assert np.array_equal(yq[:, c, :].reshape(-1), q), c
        assert np.array_equal(u32(d[:, c]), u32(dq)), c
        sums16 = bs.reshape(nsb, 4, 4).astype(np.int32)
        got16 = s16[:, c, :, 0, :].astype(np.int32) + 64 * s16[:, c, :, 1, :].view(np.int8).astype(np.int32)
        assert np.array_equal(got16, sums16), c
        assert (s16[:, c, :, 0, :] < 64).all()
        sums32 = bs.reshape(nsb, 2, 4, 2).astype(np.int32).sum(axis=3)
        got32 = s32[:, c, :, 0, :].astype(np.int32) + 64 * s32[:, c, :, 1, :].view(np.int8).astype(np.int32)
        assert np.array_equal(got32, sums32), c
    for c in range(n, 8):
        assert not d[:, c].any() and not s32[:, c].any() and not s16[:, c].any()


def test_quantize_list_columns_golden(kern, golden2):
    """the reference's own quantize_row_q8_K_ref outputs (tests/golden/golden_v2.npz) through k_quantize_cols"""
    x = np.stack([golden2["q8_K_x"], golden2["q8_K_x"]])
    k = x.shape[1]
    buf = kern.quantize_acts_columns(kern.to_device(x)).cpu().numpy()
    yq = buf[:8 * k].view(np.int8).reshape(k // 32, 8, 32)
    off = (8 * k + 127) & ~127
    d = buf[off:off + k // 256 * 32].view(np.float32).reshape(-1, 8)
    s16 = buf[off + k // 256 * 160: off + k // 256 * 416].reshape(-1, 8, 4, 2, 4)
    for c in range(2):
        assert np.array_equal(yq[:, c, :].reshape(-1), golden2["q8_K_q"])
        assert np.array_equal(u32(d[:, c]), u32(golden2["q8_K_d"]))
        got16 = s16[:, c, :, 0, :].astype(np.int32) + 64 * s16[:, c, :, 1, :].view(np.int8).astype(np.int32)
        assert np.array_equal(got16.reshape(-1), golden2["q8_K_bsums"].astype(np.int32))


def test_get_rows_index_out_of_range_is_reported(kern, oracle):
    """the reference asserts on a row index outside the table (src/ggml-cpu/ops.cpp:5147); here the row is zero-filled
    (never stale memory), the good rows are right and the library's sticky error flag is raised."""
    import b200
    import torch
    t, n_vocab, k = "q4_K", 100, 1024
    blocks = random_blocks(t, n_vocab, k, seed=5)
    rows = np.array([3, n_vocab, 7, -1, 99], dtype=np.int32)
    assert kern.device_error_flags() == 0
    got = kern.get_rows(t, dev_blocks(kern, blocks), k, torch.from_numpy(rows).to(kern.device)).cpu().numpy()
    assert kern.device_error_flags() & b200.ERR_FLAG_INDEX
    assert kern.device_error_flags() == 0            # reading clears it
    good = [0, 2, 4]
    assert np.array_equal(u32(got[good]), u32(oracle.get_rows(t, blocks, k, rows[good])))
    assert not got[[1, 3]].any()


# ---------------------------------------------------------------------------- MUL_MAT, decode (n <= 8)
def _mmvq_both(kern, monkeypatch, t, w, m, k, x):
    """run the matvec through the streaming kernel (default) and the generic kernel"""
    outs = {}
    for mode in ("stream", "generic"):
        with kern.option("mmvq_generic", 1 if mode == "generic" else 0):
            xd = kern.to_device(x)
            scratch = kern.quantize_acts(xd)
            outs[mode] = kern.mul_mat_vec_q(t, dev_blocks(kern, w), m, k, scratch, x.shape[0]).cpu().numpy()
    return outs


@pytest.mark.parametrize("t", QUANT_TYPES)
@pytest.mark.parametrize("n", [1, 3, 8])
def test_mul_mat_vec_golden(kern, oracle, golden, monkeypatch, t, n):
    """m=32, k=512 against ggml-cpu's own MUL_MAT output (generated from the unmodified reference)."""
    w = golden[f"mm_{t}_w"]
    x = golden["mm_x"][:n]
    want = golden[f"mm_{t}_y_n{n}"]
    for mode, got in _mmvq_both(kern, monkeypatch, t, w, 32, 512, x).items():
        assert oracle.nmse(want, got) <= NMSE_REF, (mode, oracle.nmse(want, got))
        assert oracle.nmse(oracle.mul_mat(t, w, 32, 512, x, mode=1), got) <= NMSE_ORACLE, mode


MMVQ_SHAPES = [(16, 256), (1, 256), (5, 512), (33, 1024), (257, 4096), (64, 11008), (48, 14336), (7, 2048), (130, 8192)]


@pytest.mark.parametrize("t", QUANT_TYPES)
@pytest.mark.parametrize("m,k", MMVQ_SHAPES)
@pytest.mark.parametrize("n", [1, 2, 4, 5, 8])
def test_mul_mat_vec_vs_oracle(kern, oracle, monkeypatch, t, m, k, n):
    w = random_blocks(t, m, k, seed=100 + m + k)
    x = np.random.default_rng(7 + n).uniform(-1, 1, size=(n, k)).astype(np.float32)
    want = oracle.mul_mat(t, w, m, k, x, mode=1)
    exact = oracle.mul_mat(t, w, m, k, x, mode=0)
    outs = _mmvq_both(kern, monkeypatch, t, w, m, k, x)
    for mode, got in outs.items():
        assert got.shape == (n, m)
        assert oracle.nmse(want, got) <= NMSE_ORACLE, (mode, oracle.nmse(want, got))
        if m * n >= 64:   # q8_1 rounding noise is a per-element statistic: meaningless on a handful of outputs
            assert oracle.nmse(exact, got) <= NMSE_REF, (mode, oracle.nmse(exact, got))


@pytest.mark.parametrize("t", ["q4_0", "q8_0"])
def test_mul_mat_vec_odd_block_count(kern, oracle, monkeypatch, t):
    """k = 32 * odd: no block pairs -> generic kernel; also k = 32 (a single block)."""
    for m, k in [(16, 32), (9, 96), (20, 32 * 37)]:
        w = random_blocks(t, m, k, seed=k)
        x = np.random.default_rng(k).uniform(-1, 1, size=(2, k)).astype(np.float32)
        for mode, got in _mmvq_both(kern, monkeypatch, t, w, m, k, x).items():
            assert oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=1), got) <= NMSE_ORACLE, (mode, m, k)


@pytest.mark.parametrize("t", QUANT_TYPES)
def test_mul_mat_vec_row_stride_and_offset(kern, oracle, t):
    """weights that are a strided view (row stride 2x) starting at a 2-byte-aligned offset"""
    import torch
    m, k = 37, 1024
    rb = k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]
    w = random_blocks(t, m, k, seed=77)
    for lead in (0, 2 * BLOCK_BYTES[t]):
        stride = 2 * rb
        buf = np.zeros(lead + m * stride + 64, dtype=np.uint8)
        for r in range(m):
            buf[lead + r * stride: lead + r * stride + rb] = w[r]
        x = np.random.default_rng(5).uniform(-1, 1, size=(3, k)).astype(np.float32)
        xd = kern.to_device(x)
        scratch = kern.quantize_acts(xd)
        wd = kern.to_device(buf)
        got = kern.mul_mat_vec_q(t, wd[lead:], m, k, scratch, 3, w_row_stride=stride).cpu().numpy()
        assert oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=1), got) <= NMSE_ORACLE, lead


@pytest.mark.parametrize("t", QUANT_TYPES)
def test_mul_mat_4d_broadcast(kern, oracle, t):
    """ggml broadcast: a [k, m, 2, 1] x b [k, n, 4, 3] -> [m, n, 4, 3] (bs={2,1}, nr={2,3})"""
    m, k, n = 16, 256, 3
    ne02, ne03, ne12, ne13 = 2, 1, 4, 3
    rb = k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]
    w = random_blocks(t, m * ne02 * ne03, k, seed=9)
    x = np.random.default_rng(10).uniform(-1, 1, size=(ne13, ne12, n, k)).astype(np.float32)
    ne_a = [k, m, ne02, ne03]
    nb_a = [BLOCK_BYTES[t], rb, rb * m, rb * m * ne02]
    ne_b = [k, n, ne12, ne13]
    nb_b = [4, 4 * k, 4 * k * n, 4 * k * n * ne12]
    got = kern.mul_mat_4d(t, dev_blocks(kern, w), ne_a, nb_a, kern.to_device(x), ne_b, nb_b).cpu().numpy()
    wv = w.reshape(ne03, ne02, m, rb)
    for i13 in range(ne13):
        for i12 in range(ne12):
            wch = wv[i13 // (ne13 // ne03), i12 // (ne12 // ne02)]
            want = oracle.mul_mat(t, wch, m, k, x[i13, i12], mode=1)
            assert oracle.nmse(want, got[i13, i12]) <= NMSE_ORACLE, (i13, i12)


@pytest.mark.parametrize("t", QUANT_TYPES)
def test_mul_mat_4d_permuted_b(kern, oracle, t):
    """b permuted {0,2,1,3}: memory layout [ne13][n][ne12][k], logical [k, n, ne12, ne13]"""
    m, k, n, ne12 = 16, 256, 8, 3
    rb = k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]
    w = random_blocks(t, m, k, seed=19)
    xm = np.random.default_rng(20).uniform(-1, 1, size=(n, ne12, k)).astype(np.float32)   # memory order
    ne_a = [k, m, 1, 1]
    nb_a = [BLOCK_BYTES[t], rb, rb * m, rb * m]
    ne_b = [k, n, ne12, 1]
    nb_b = [4, 4 * k * ne12, 4 * k, 4 * k * ne12 * n]
    got = kern.mul_mat_4d(t, dev_blocks(kern, w), ne_a, nb_a, kern.to_device(xm), ne_b, nb_b).cpu().numpy()
    for i12 in range(ne12):
        want = oracle.mul_mat(t, w, m, k, np.ascontiguousarray(xm[:, i12]), mode=1)
        assert oracle.nmse(want, got[0, i12]) <= NMSE_ORACLE, i12


# ---------------------------------------------------------------------------- MUL_MAT n = 1: the fused decode kernel
FUSED_TYPES = ["q4_K", "q5_K"]
FUSED_SHAPES = [(1, 1024), (2, 1024), (3, 4096), (37, 1280), (600, 4096), (1000, 2048), (129, 8192), (300, 11008),
                (77, 14336), (40, 28672), (20, 32768), (5000, 4096)]


@pytest.mark.parametrize("t", FUSED_TYPES)
@pytest.mark.parametrize("m,k", FUSED_SHAPES)
def test_mul_mat_fused_decode_vs_oracle(kern, oracle, monkeypatch, t, m, k):
    """b200_mul_mat with n = 1 takes the fused quantize + mat-vec kernel (mmvq_fused.cuh): same q8_1
    arithmetic as the two-kernel path -> NMSE 1e-9 to oracle mode 1, and row r is bit-identical to the
    two-kernel path's partial-order-independent value only up to f32 summation order, so the cross
    check between the two paths is NMSE as well."""
    w = random_blocks(t, m, k, seed=3 * m + k)
    x = np.random.default_rng(m + k).uniform(-1, 1, size=(1, k)).astype(np.float32)
    wd, xd = dev_blocks(kern, w), kern.to_device(x)
    want = oracle.mul_mat(t, w, m, k, x, mode=1)
    for const in (False, True):
        got = kern.mul_mat(t, wd, m, k, xd, const_weights=const).cpu().numpy().reshape(1, m)
        assert np.isfinite(got).all()
        assert oracle.nmse(want, got) <= NMSE_ORACLE, (const, oracle.nmse(want, got))
    with kern.option("mmvq_no_fused", 1):
        two = kern.mul_mat(t, wd, m, k, xd).cpu().numpy().reshape(1, m)
    assert oracle.nmse(two, got) <= NMSE_ORACLE


@pytest.mark.parametrize("t", FUSED_TYPES)
def test_mul_mat_fused_decode_adversarial_scales_and_activations(kern, oracle, t):
    """edge scale fields (0, subnormal, max, sc = m = 63) and activations with zero blocks, a single
    outlier and +-x.5 ties (the q8_1 rounding contract)."""
    m, k = 64, 4096
    w = adversarial_blocks(t, m, k, seed=11)
    x = np.random.default_rng(12).standard_normal(size=(1, k)).astype(np.float32)
    x[0, :64] = 0.0
    x[0, 100] = 1000.0
    x[0, 128:160] = (np.arange(32) - 16 + 0.5) * (127.0 / 15.5) / 127.0
    got = kern.mul_mat(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy().reshape(1, m)
    want = oracle.mul_mat(t, w, m, k, x, mode=1)
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin)
    assert oracle.nmse(want[fin], got[fin]) <= NMSE_ORACLE


@pytest.mark.parametrize("t", FUSED_TYPES)
def test_mul_mat_fused_decode_row_stride_and_row_independence(kern, oracle, t):
    """strided weight view (row stride 2x) and the row-split property: a row's value does not depend
    on which other rows the launch covers (bit-identical sub-matrix results)."""
    m, k = 211, 4096
    rb = k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]
    w = random_blocks(t, m, k, seed=5)
    x = np.random.default_rng(6).uniform(-1, 1, size=(1, k)).astype(np.float32)
    xd = kern.to_device(x)
    full = kern.mul_mat(t, dev_blocks(kern, w), m, k, xd).cpu().numpy().reshape(m)
    buf = np.zeros(m * 2 * rb + 64, dtype=np.uint8)
    for r in range(m):
        buf[r * 2 * rb: r * 2 * rb + rb] = w[r]
    strided = kern.mul_mat(t, kern.to_device(buf), m, k, xd, w_row_stride=2 * rb).cpu().numpy().reshape(m)
    assert np.array_equal(u32(full), u32(strided))
    lo, hi = 64, 173
    part = kern.mul_mat(t, dev_blocks(kern, w[lo:hi]), hi - lo, k, xd).cpu().numpy().reshape(hi - lo)
    assert np.array_equal(u32(part), u32(full[lo:hi]))


def test_mul_mat_batch_matches_single_calls(kern, oracle):
    import b200
    import torch
    mats = [("q4_K", 300, 4096), ("q5_K", 128, 2048), ("q6_K", 64, 1024), ("q4_0", 96, 512), ("q4_K", 17, 11008)]
    ws = [dev_blocks(kern, random_blocks(t, m, k, seed=i)) for i, (t, m, k) in enumerate(mats)]
    xs = [kern.to_device(np.random.default_rng(i).uniform(-1, 1, size=(1, k)).astype(np.float32)) for i, (t, m, k) in enumerate(mats)]
    ys = [torch.zeros((1, m), dtype=torch.float32, device=kern.device) for (t, m, k) in mats]
    nodes = (b200.MulMatNode * len(mats))(*[
        b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), 1, y.data_ptr(), const_weights=True)
        for (t, m, k), w, x, y in zip(mats, ws, xs, ys)])
    kern.mul_mat_batch(nodes)
    for (t, m, k), w, x, y in zip(mats, ws, xs, ys):
        single = kern.mul_mat(t, w, m, k, x)
        if t not in FUSED_TYPES:      # list: persistent kernel; single call: stream kernel -- same q8_1 arithmetic, another f32 summation order
            assert oracle.nmse(single.cpu().numpy().reshape(1, -1), y.cpu().numpy().reshape(1, -1)) <= NMSE_ORACLE
            continue
        assert np.array_equal(u32(single.cpu().numpy().reshape(-1)), u32(y.cpu().numpy().reshape(-1))), (t, m, k)


@pytest.mark.parametrize("n", [2, 3, 4, 8])
def test_mul_mat_batch_few_columns(kern, oracle, n):
    """decode at batch n: nodes with 2..8 columns go to the MMA list kernel (mmvq_cols.cuh: all five formats; Q6_K when its row
    is a whole number of 8-super-block chunks); what does not qualify rides the persistent launch as one entry per column (same
    weights, L2-resident for the later columns), where each column equals the same column computed alone (n = 1 list) bit
    for bit.  All within the NMSE bound of the oracle.  Column stride of x is padded (nb_b[1] > 4 k)."""
    import b200
    import torch
    mats = [("q4_K", 300, 4096), ("q6_K", 130, 1024), ("q8_0", 96, 2048), ("q5_K", 4096, 1024), ("q4_0", 64, 4096),
            ("q4_K", 17, 256), ("q5_K", 33, 11008), ("q4_K", 4096, 14336), ("q6_K", 200, 4096), ("q6_K", 49, 14336), ("q6_K", 1024, 2048),
            ("q6_K", 77, 11008), ("q6_K", 33, 2304), ("q6_K", 1, 256), ("q8_0", 40, 4128),     # Q6_K rows at 2-byte alignment; k % 256 != 0
            ("q4_0", 33, 256), ("q8_0", 1, 11008), ("q4_0", 4097, 1024), ("q8_0", 300, 28672), ("q4_0", 31, 11008), ("q6_K", 4099, 768)]
    c0 = kern.launch_count()
    for i, (t, m, k) in enumerate(mats):
        w_np = random_blocks(t, m, k, seed=40 + i)
        w = dev_blocks(kern, w_np)
        pad = 32
        x_np = np.random.default_rng(50 + i).uniform(-1, 1, size=(n, k + pad)).astype(np.float32)
        x = kern.to_device(x_np)
        y = torch.zeros((n, m), dtype=torch.float32, device=kern.device)
        nodes = (b200.MulMatNode * 1)(b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), n, y.data_ptr(), b_row_stride=4 * (k + pad)))
        kern.mul_mat_batch(nodes)
        got = y.cpu().numpy()
        # route 2 = the MMA list kernel (every format when k % 256 == 0; Q6_K: chunks of 8 super-blocks must be 16-byte multiples)
        on_mma_kernel = b200.batch_plan(kern.lib, nodes)[0][2] == 2
        if not any(v.startswith("GGML_B200_COLS") or v == "GGML_B200_BATCH_COLS" for v in os.environ):      # default routing
            assert on_mma_kernel or (t == "q8_0" and k % 256), (t, m, k)
        # the MMA list kernel takes its activations through Q8_K like ggml-cpu (oracle mode 2), the others through q8_1
        want = oracle.mul_mat(t, w_np, m, k, np.ascontiguousarray(x_np[:, :k]), mode=2 if on_mma_kernel else 1)
        assert oracle.nmse(want, got) <= NMSE_ORACLE, (t, m, k, n)
        if on_mma_kernel:
            # the MMA list kernel (mmvq_cols.cuh): a column's value depends on that column alone -- the first two
            # columns computed as an n = 2 node give the same bits; and the result is reproducible
            y2 = torch.zeros((2, m), dtype=torch.float32, device=kern.device)
            kern.mul_mat_batch((b200.MulMatNode * 1)(b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), 2, y2.data_ptr(), b_row_stride=4 * (k + pad))))
            assert np.array_equal(u32(y2.cpu().numpy().reshape(-1)), u32(got[:2].reshape(-1))), (t, m, k, n)
            y.zero_()
            kern.mul_mat_batch(nodes)
            assert np.array_equal(u32(y.cpu().numpy().reshape(-1)), u32(got.reshape(-1))), (t, m, k, n)
        elif n <= 4:
            for c in range(n):
                y1 = torch.zeros((1, m), dtype=torch.float32, device=kern.device)
                xc = kern.to_device(np.ascontiguousarray(x_np[c:c + 1, :k]))
                one = (b200.MulMatNode * 2)(b200.mul_mat_node(t, w.data_ptr(), m, k, xc.data_ptr(), 1, y1.data_ptr()),
                                            b200.mul_mat_node(t, w.data_ptr(), m, k, xc.data_ptr(), 1, y1.data_ptr()))
                with kern.option("batch_single", 1):
                    kern.mul_mat_batch((b200.MulMatNode * 1)(one[0]))
                assert np.array_equal(u32(y1.cpu().numpy().reshape(-1)), u32(got[c])), (t, m, k, n, c)
    assert kern.launch_count() > c0


def test_mul_mat_batch_few_columns_padded_row_stride(kern, oracle):
    """weight rows padded to a 16-byte multiple, read with the padded stride; Q6_K with a super-block count that is not a multiple
    of 8 (its last chunk is not a multiple of 16 bytes) rides the ring family for rows at 2-byte alignment"""
    import b200
    import torch
    for t, m, k, mode in (("q6_K", 70, 2304, 2), ("q6_K", 70, 2048, 2), ("q4_K", 70, 2304, 2), ("q5_K", 33, 768, 2)):
        w_np = random_blocks(t, m, k, seed=91)
        rb = w_np.shape[1]
        stride = (rb + 15) // 16 * 16 + 32
        wp = np.zeros((m, stride), dtype=np.uint8)
        wp[:, :rb] = w_np
        w = kern.to_device(wp.reshape(-1))
        x_np = np.random.default_rng(92).uniform(-1, 1, size=(4, k)).astype(np.float32)
        x = kern.to_device(x_np)
        y = torch.zeros((4, m), dtype=torch.float32, device=kern.device)
        nodes = (b200.MulMatNode * 1)(b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), 4, y.data_ptr(), a_row_stride=stride))
        route = b200.batch_plan(kern.lib, nodes)[0][2]
        if not any(v.startswith("GGML_B200_COLS") for v in os.environ):
            assert route == 2
        kern.mul_mat_batch(nodes)
        assert oracle.nmse(oracle.mul_mat(t, w_np, m, k, x_np, mode=mode if route == 2 else 1), y.cpu().numpy()) <= NMSE_ORACLE, (t, m, k)


def test_mul_mat_batch_prefill_nodes_share_lanes(kern, oracle):
    """independent batched (n > 8) nodes of a list run side by side on up to three streams; a node that reads another's
    output waits for it.  Every output equals the node computed alone, bit for bit -- eagerly and replayed from a CUDA graph
    (the side streams are captured as parallel branches)."""
    import b200
    import torch
    n = 48
    specs = [("q4_K", 1024, 1024, "x"), ("q6_K", 512, 1024, "x"), ("q8_0", 300, 1024, "x"), ("q4_K", 640, 1024, 0),    # node 3 reads node 0
             ("q5_K", 256, 512, "x2"), ("q4_0", 200, 640, 3)]                                                         # node 5 reads node 3
    rng = np.random.default_rng(77)
    xs = {"x": kern.to_device(rng.uniform(-1, 1, size=(n, 1024)).astype(np.float32)),
          "x2": kern.to_device(rng.uniform(-1, 1, size=(n, 512)).astype(np.float32))}
    ws = [dev_blocks(kern, random_blocks(t, m, k, seed=80 + i)) for i, (t, m, k, _) in enumerate(specs)]
    ys = [torch.zeros((n, m), dtype=torch.float32, device=kern.device) for (_, m, _, _) in specs]
    src = lambda i: xs[specs[i][3]] if isinstance(specs[i][3], str) else ys[specs[i][3]]
    nodes = (b200.MulMatNode * len(specs))(*[b200.mul_mat_node(t, ws[i].data_ptr(), m, k, src(i).data_ptr(), n, ys[i].data_ptr())
                                             for i, (t, m, k, _) in enumerate(specs)])
    wsz = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(nodes, len(specs))) + 256, dtype=torch.uint8, device=kern.device)
    kern.mul_mat_batch(nodes, ws=wsz)
    torch.cuda.synchronize()
    want = []
    for i, (t, m, k, _) in enumerate(specs):
        xin = src(i) if isinstance(specs[i][3], str) else want[specs[i][3]]
        want.append(kern.mul_mat(t, ws[i], m, k, xin.contiguous()).reshape(n, m).clone())
    for i in range(len(specs)):
        assert torch.equal(ys[i].view(torch.int32), want[i].view(torch.int32)), i
    # replay from a CUDA graph
    for y in ys:
        y.zero_()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        kern.mul_mat_batch(nodes, ws=wsz)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    for i in range(len(specs)):
        assert torch.equal(ys[i].view(torch.int32), want[i].view(torch.int32)), ("graph", i)


def test_mul_mat_batch_few_columns_long_mixed_list(kern, oracle):
    """more 2..8-column nodes than one launch of the MMA list kernel takes (384 per family): all five formats interleaved,
    groups of three sharing their activations (q / k / v), row counts that are not multiples of the 32-row tile, k up to
    28672 (14 chunks of the ring).  Sampled nodes against the oracle, all of them finite and reproducible."""
    import b200
    import torch
    n = 5
    rng = np.random.default_rng(123)
    types = ("q4_K", "q6_K", "q5_K", "q4_0", "q8_0", "q6_K")
    shapes = [(33, 2048), (70, 4096), (17, 2048), (129, 2048), (40, 1280)]     # k = 1280: Q6_K rows at 2-byte alignment (family 4)
    nodes, keep = [], []
    xs = {}
    for i in range(1000):
        t = types[(i % 3) + 3 * ((i // 3) % 2)]
        m, k = (24, 28672) if i == 500 else shapes[(i // 3) % len(shapes)]
        g = i // 3                                   # three consecutive nodes read the same x
        if (g, k) not in xs:
            xs[(g, k)] = kern.to_device(rng.uniform(-1, 1, size=(n, k)).astype(np.float32))
        x = xs[(g, k)]
        w_np = random_blocks(t, m, k, seed=1000 + i) if i % 97 == 0 or i == 500 else None
        w = dev_blocks(kern, w_np) if w_np is not None else kern.to_device(random_blocks(t, m, k, seed=7).reshape(-1))
        y = torch.zeros((n, m), dtype=torch.float32, device=kern.device)
        nodes.append(b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), n, y.data_ptr()))
        keep.append((t, m, k, w, w_np, x, y))
    arr = (b200.MulMatNode * len(nodes))(*nodes)
    wsz = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(arr, len(nodes))) + 256, dtype=torch.uint8, device=kern.device)
    kern.mul_mat_batch(arr, ws=wsz)
    routes = [r for (_, _, r) in b200.batch_plan(kern.lib, arr)]
    first = [y.clone() for (_, _, _, _, _, _, y) in keep]
    for (t, m, k, w, w_np, x, y), route in zip(keep, routes):
        assert torch.isfinite(y).all()
        if w_np is not None:
            want = oracle.mul_mat(t, w_np, m, k, x.cpu().numpy(), mode=2 if route == 2 else 1)
            assert oracle.nmse(want, y.cpu().numpy()) <= NMSE_ORACLE, (t, m, k)
    for (_, _, _, _, _, _, y) in keep:
        y.zero_()
    kern.mul_mat_batch(arr, ws=wsz)
    for a, (_, _, _, _, _, _, y) in zip(first, keep):
        assert torch.equal(a.view(torch.int32), y.view(torch.int32))


def _batch_nodes(kern, mats, seed0=0, shared_x=False):
    import b200
    import torch
    ws_np = [random_blocks(t, m, k, seed=seed0 + i) for i, (t, m, k) in enumerate(mats)]
    ws = [dev_blocks(kern, w) for w in ws_np]
    xs_np, cache = [], {}
    for i, (t, m, k) in enumerate(mats):
        if shared_x and k in cache:
            xs_np.append(cache[k])
        else:
            cache[k] = np.random.default_rng(seed0 + 100 + i).uniform(-1, 1, size=(1, k)).astype(np.float32)
            xs_np.append(cache[k])
    xdev = {}
    xs = []
    for x in xs_np:
        if id(x) not in xdev:
            xdev[id(x)] = kern.to_device(x)
        xs.append(xdev[id(x)])
    ys = [torch.zeros((1, m), dtype=torch.float32, device=kern.device) for (t, m, k) in mats]
    nodes = (b200.MulMatNode * len(mats))(*[
        b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), 1, y.data_ptr(), const_weights=True)
        for (t, m, k), w, x, y in zip(mats, ws, xs, ys)])
    return nodes, ws_np, ws, xs_np, xs, ys


@pytest.mark.parametrize("shared_x", [False, True])
def test_mul_mat_persistent_batch_vs_oracle(kern, oracle, shared_x):
    """the persistent list kernel (mmvq_batch.cuh): mixed formats, K from 1024 to 32768 (W = 1..8 warps
    per row), odd and tiny m, m below / far above the SM count, shared and distinct activations.  Every
    node: NMSE 1e-9 to the oracle's q8_1 arithmetic and bit-identical to the single-node call."""
    mats = [("q4_K", 300, 4096), ("q5_K", 128, 2048), ("q4_K", 17, 11008), ("q4_K", 1, 1024), ("q5_K", 3, 4096),
            ("q4_K", 4096, 4096), ("q5_K", 1111, 11008), ("q4_K", 40, 28672), ("q5_K", 21, 32768), ("q4_K", 2, 1280),
            ("q4_K", 999, 4096), ("q4_K", 5000, 1024), ("q5_K", 77, 14336), ("q4_K", 148 * 2 + 1, 8192),
            ("q6_K", 4096, 4096), ("q6_K", 333, 11008), ("q6_K", 3, 1024), ("q6_K", 50, 14336), ("q6_K", 1000, 2304),
            ("q4_0", 4096, 4096), ("q8_0", 300, 4096), ("q4_0", 77, 11008), ("q8_0", 50, 1024), ("q8_0", 1001, 2304),
            ("q4_0", 5, 32768), ("q8_0", 33, 14336)]
    nodes, ws_np, ws, xs_np, xs, ys = _batch_nodes(kern, mats, seed0=40, shared_x=shared_x)
    c0 = kern.launch_count()
    kern.mul_mat_batch(nodes)
    assert kern.launch_count() - c0 == 6, "independent decode nodes ride in one (quantize, persistent) launch pair per format family"
    for (t, m, k), w_np, w, x_np, x, y in zip(mats, ws_np, ws, xs_np, xs, ys):
        got = y.cpu().numpy().reshape(1, m)
        want = oracle.mul_mat(t, w_np, m, k, x_np, mode=1)
        assert oracle.nmse(want, got) <= NMSE_ORACLE, (t, m, k, oracle.nmse(want, got))
        if t in FUSED_TYPES:
            single = kern.mul_mat(t, w, m, k, x).cpu().numpy().reshape(-1)
        else:
            # formats whose single-node call takes another kernel: the row values must still not depend on the
            # list -- the same node alone in a persistent launch gives the same bits
            import b200
            import torch
            y1 = torch.zeros((1, m), dtype=torch.float32, device=kern.device)
            one = (b200.MulMatNode * 1)(b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), 1, y1.data_ptr()))
            with kern.option("batch_single", 1):
                kern.mul_mat_batch(one)
            single = y1.cpu().numpy().reshape(-1)
        assert np.array_equal(u32(single), u32(got.reshape(-1))), (t, m, k)


@pytest.mark.parametrize("t", ["q6_K", "q4_0", "q8_0"])
def test_mul_mat_persistent_unaligned_formats_adversarial_and_strided(kern, oracle, t):
    """the 2-byte-aligned formats through the persistent kernel: edge scales (0, subnormal, max d; Q6_K scales
    +-127/-128), rows that are not 16-byte aligned (Q6_K k = 2304: 1890-byte rows; a strided view whose rows start
    at every 2-byte phase), zero / outlier / tie activations."""
    import b200
    import torch
    m, k = 96, 2304
    rb = k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]
    w_np = adversarial_blocks(t, m, k, seed=21)
    buf = np.zeros(m * (rb + 6) + 64, dtype=np.uint8)
    for r in range(m):
        buf[r * (rb + 6): r * (rb + 6) + rb] = w_np[r]
    x_np = np.random.default_rng(22).standard_normal(size=(1, k)).astype(np.float32)
    x_np[0, :64] = 0.0
    x_np[0, 100] = 1000.0
    x_np[0, 128:160] = (np.arange(32) - 16 + 0.5) * (127.0 / 15.5) / 127.0
    w, ws, x = dev_blocks(kern, w_np), kern.to_device(buf), kern.to_device(x_np)
    y = torch.zeros((2, m), dtype=torch.float32, device=kern.device)
    nodes = (b200.MulMatNode * 2)(b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), 1, y[0].data_ptr()),
                                  b200.mul_mat_node(t, ws.data_ptr(), m, k, x.data_ptr(), 1, y[1].data_ptr(), a_row_stride=rb + 6))
    kern.mul_mat_batch(nodes)
    got = y.cpu().numpy()
    want = oracle.mul_mat(t, w_np, m, k, x_np, mode=1)
    fin = np.isfinite(want[0])
    assert np.array_equal(np.isfinite(got[0]), fin)
    assert oracle.nmse(want[0][fin], got[0][fin]) <= NMSE_ORACLE
    assert np.array_equal(u32(got[0][fin]), u32(got[1][fin]))


def test_mul_mat_persistent_batch_many_nodes_and_strided_rows(kern, oracle):
    """more nodes than one launch holds (448) and a strided weight view inside the list."""
    import b200
    import torch
    t, m, k = "q4_K", 64, 1024
    rb = k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]
    w_np = random_blocks(t, m, k, seed=9)
    buf = np.zeros(m * 2 * rb + 64, dtype=np.uint8)
    for r in range(m):
        buf[r * 2 * rb: r * 2 * rb + rb] = w_np[r]
    w, wstr = dev_blocks(kern, w_np), kern.to_device(buf)
    count = 460
    x_np = np.random.default_rng(3).uniform(-1, 1, size=(count, k)).astype(np.float32)
    x = kern.to_device(x_np)
    y = torch.zeros((count, m), dtype=torch.float32, device=kern.device)
    nodes = (b200.MulMatNode * count)(*[
        b200.mul_mat_node(t, (wstr if i % 3 == 1 else w).data_ptr(), m, k, x[i].data_ptr(), 1, y[i].data_ptr(),
                          a_row_stride=2 * rb if i % 3 == 1 else None, const_weights=(i % 2 == 0))
        for i in range(count)])
    c0 = kern.launch_count()
    kern.mul_mat_batch(nodes)
    assert kern.launch_count() - c0 == 4        # 448 + 12 nodes: two (quantize, persistent) pairs
    got = y.cpu().numpy()
    want = oracle.mul_mat(t, w_np, m, k, x_np, mode=1)
    assert oracle.nmse(want, got) <= NMSE_ORACLE


def test_mul_mat_persistent_batch_chains_dependencies(kern, oracle):
    """a node that reads (or overwrites) what an earlier node of the list writes: chain y1 = W1 x, y2 = W2 y1, then a
    node that overwrites y1.  Default: all three ride ONE persistent launch (per-node completion counters, the
    dependent node's activations quantized inside the launch); option no_chain = 1: the list is cut at every hazard
    (three launch groups).  Same bits either way, and the same as three separate calls."""
    import b200
    import torch
    w1_np, w2_np, w3_np = random_blocks("q4_K", 4096, 1024, seed=1), random_blocks("q5_K", 200, 4096, seed=2), random_blocks("q4_K", 4096, 2048, seed=3)
    w1, w2, w3 = dev_blocks(kern, w1_np), dev_blocks(kern, w2_np), dev_blocks(kern, w3_np)
    x_np = np.random.default_rng(8).uniform(-1, 1, size=(1, 1024)).astype(np.float32)
    x3_np = np.random.default_rng(9).uniform(-1, 1, size=(1, 2048)).astype(np.float32)
    x, x3 = kern.to_device(x_np), kern.to_device(x3_np)
    y1 = torch.zeros((1, 4096), dtype=torch.float32, device=kern.device)
    y2 = torch.zeros((1, 200), dtype=torch.float32, device=kern.device)
    nodes = (b200.MulMatNode * 3)(
        b200.mul_mat_node("q4_K", w1.data_ptr(), 4096, 1024, x.data_ptr(), 1, y1.data_ptr()),
        b200.mul_mat_node("q5_K", w2.data_ptr(), 200, 4096, y1.data_ptr(), 1, y2.data_ptr()),
        b200.mul_mat_node("q4_K", w3.data_ptr(), 4096, 2048, x3.data_ptr(), 1, y1.data_ptr()))
    r1 = kern.mul_mat("q4_K", w1, 4096, 1024, x)
    r2 = kern.mul_mat("q5_K", w2, 200, 4096, r1)
    r3 = kern.mul_mat("q4_K", w3, 4096, 2048, x3)
    for no_chain, launches in ((0, 2), (1, 4)):
        y1.zero_(); y2.zero_()
        with kern.option("no_chain", no_chain):
            c0 = kern.launch_count()
            kern.mul_mat_batch(nodes)
            # chained: one (quantize, persistent) pair.  cut: three groups of one node -- the two small ones take the
            # per-node kernel (1 launch each), the 4.7 MB one the (quantize, persistent) pair
            assert kern.launch_count() - c0 == launches, (no_chain, kern.launch_count() - c0)
        assert np.array_equal(u32(y2.cpu().numpy().reshape(-1)), u32(r2.cpu().numpy().reshape(-1))), no_chain
        assert np.array_equal(u32(y1.cpu().numpy().reshape(-1)), u32(r3.cpu().numpy().reshape(-1))), no_chain
    assert oracle.nmse(oracle.mul_mat("q5_K", w2_np, 200, 4096, r1.cpu().numpy().reshape(1, 4096), mode=1), y2.cpu().numpy()) <= NMSE_ORACLE


def _decoder_chain(kern, types, hidden, ffn, layers, vocab, seed):
    """The dependency structure of a decoder pass on small shapes: per layer q/k/v read h, wo reads q (stand-in for the
    attention output), gate/up read wo, down reads up; the next layer reads down; the output matrix reads the last
    down.  Returns (nodes, weights (numpy, device), buffers, order) for b200_mul_mat_batch."""
    import b200
    import torch
    tq, tk, tv, to, tg, tu, td, tout = types
    dev = kern.device
    h0_np = np.random.default_rng(seed).uniform(-1, 1, size=(1, hidden)).astype(np.float32)
    h0 = kern.to_device(h0_np)
    specs, bufs = [], {"h0": h0}
    cur = "h0"
    for l in range(layers):
        for nm, t in (("q", tq), ("k", tk), ("v", tv)):
            specs.append((t, hidden, hidden, cur, f"{nm}{l}"))
        specs.append((to, hidden, hidden, f"q{l}", f"o{l}"))
        specs.append((tg, ffn, hidden, f"o{l}", f"g{l}"))
        specs.append((tu, ffn, hidden, f"o{l}", f"u{l}"))
        specs.append((td, hidden, ffn, f"u{l}", f"d{l}"))
        cur = f"d{l}"
    specs.append((tout, vocab, hidden, cur, "logits"))
    # random blocks amplify (|y| ~ 30 |x| per node): without the normalisations of a real model the chain would overflow
    # after a few layers, so every node's scale fields are calibrated to unit output RMS on the way (a stand-in for them)
    from oracle_lib import Oracle, scale_blocks
    orc = Oracle()
    ws_np, ws, vals = [], [], {"h0": h0_np}
    for i, (t, m, k, src, dst) in enumerate(specs):
        w = random_blocks(t, m, k, seed=seed + 13 * i)
        y = orc.mul_mat(t, w, m, k, vals[src].reshape(1, k), mode=1)
        scale_blocks(t, w, 1.0 / max(float(np.sqrt(np.mean(y.astype(np.float64) ** 2))), 1e-30))
        vals[dst] = orc.mul_mat(t, w, m, k, vals[src].reshape(1, k), mode=1)
        ws_np.append(w); ws.append(dev_blocks(kern, w))
        bufs[dst] = torch.zeros((1, m), dtype=torch.float32, device=dev)
    nodes = (b200.MulMatNode * len(specs))(*[
        b200.mul_mat_node(t, w.data_ptr(), m, k, bufs[src].data_ptr(), 1, bufs[dst].data_ptr(), const_weights=True)
        for (t, m, k, src, dst), w in zip(specs, ws)])
    return nodes, specs, ws_np, ws, bufs, h0_np


CHAIN_MIXES = {
    "q4_K": ("q4_K",) * 8,
    "q4_K+q5_K": ("q4_K", "q5_K", "q4_K", "q5_K", "q4_K", "q4_K", "q5_K", "q4_K"),
    "q6_K": ("q6_K",) * 8,
    "q4_0+q8_0": ("q4_0", "q4_0", "q8_0", "q4_0", "q8_0", "q8_0", "q4_0", "q8_0"),
    "q4_K_M-like": ("q4_K", "q4_K", "q6_K", "q4_K", "q4_K", "q4_K", "q6_K", "q6_K"),
}


@pytest.mark.parametrize("mix", list(CHAIN_MIXES.keys()))
def test_mul_mat_chain_decoder_layers(kern, oracle, mix):
    """A whole (small) decoder pass as ONE list with its real dependency structure.  Chained inside the persistent
    launch(es) it must give, bit for bit, what the cut list (option no_chain) gives and what node-by-node calls
    give; replayed from a CUDA graph 20 times the bits never change (a missed dependency would show as a race);
    every node is within NMSE 1e-9 of the oracle fed with the node's actual input."""
    import torch
    hidden, ffn, layers, vocab = 1024, 2816, 3, 3000
    nodes, specs, ws_np, ws, bufs, h0_np = _decoder_chain(kern, CHAIN_MIXES[mix], hidden, ffn, layers, vocab, seed=500)
    names = [dst for (_, _, _, _, dst) in specs]

    def snapshot():
        return {nm: bufs[nm].cpu().numpy().copy() for nm in names}

    def clear():
        for nm in names:
            bufs[nm].zero_()

    with kern.option("no_chain", 1):
        clear()
        c0 = kern.launch_count()
        kern.mul_mat_batch(nodes)
        cut_launches = kern.launch_count() - c0
    cut = snapshot()
    clear()
    c0 = kern.launch_count()
    kern.mul_mat_batch(nodes)
    chained_launches = kern.launch_count() - c0
    got = snapshot()
    one_family = mix in ("q4_K", "q4_K+q5_K", "q6_K", "q4_0+q8_0")
    if one_family:
        assert chained_launches == 2, chained_launches          # one (quantize, persistent) pair for the whole pass
    assert chained_launches < cut_launches, (chained_launches, cut_launches)
    for nm in names:
        assert np.array_equal(u32(got[nm]), u32(cut[nm])), (mix, nm)
    # node by node, each fed with the chained run's own intermediate values
    inputs = {"h0": h0_np, **got}
    for (t, m, k, src, dst), w_np in zip(specs, ws_np):
        want = oracle.mul_mat(t, w_np, m, k, inputs[src].reshape(1, k), mode=1)
        assert oracle.nmse(want, got[dst]) <= NMSE_ORACLE, (mix, dst)
    # replay: same bits every time
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        kern.mul_mat_batch(nodes)
    for rep in range(20):
        if rep % 5 == 0:
            clear()
        g.replay()
        torch.cuda.synchronize()
        now = snapshot()
        for nm in names:
            assert np.array_equal(u32(now[nm]), u32(got[nm])), (mix, rep, nm)


def test_mul_mat_chain_write_after_read_and_reused_buffers(kern, oracle):
    """hazards other than read-after-write inside one launch: a residual-style buffer that is read, overwritten and read
    again (A = W1 x; B = W2 A; A = W3 B; C = W4 A), and two nodes writing one buffer in turn.  Same bits as the cut list."""
    import b200
    import torch
    k = 1024
    ws_np = [random_blocks("q4_K", k, k, seed=70 + i) for i in range(5)]
    ws = [dev_blocks(kern, w) for w in ws_np]
    x = kern.to_device(np.random.default_rng(71).uniform(-1, 1, size=(1, k)).astype(np.float32))
    A, B, C = (torch.zeros((1, k), dtype=torch.float32, device=kern.device) for _ in range(3))
    plan = [(0, x, A), (1, A, B), (2, B, A), (3, A, C), (4, x, C), (0, C, B)]
    nodes = (b200.MulMatNode * len(plan))(*[b200.mul_mat_node("q4_K", ws[wi].data_ptr(), k, k, src.data_ptr(), 1, dst.data_ptr()) for wi, src, dst in plan])
    outs = {}
    for no_chain in (1, 0):
        for t_ in (A, B, C):
            t_.zero_()
        with kern.option("no_chain", no_chain):
            c0 = kern.launch_count()
            kern.mul_mat_batch(nodes)
            n_launch = kern.launch_count() - c0
        outs[no_chain] = [t_.cpu().numpy().copy() for t_ in (A, B, C)] + [n_launch]
    assert outs[0][3] == 2 and outs[1][3] > 2, (outs[0][3], outs[1][3])
    for a, b in zip(outs[0][:3], outs[1][:3]):
        assert np.array_equal(u32(a), u32(b))
    # and against a step-by-step evaluation
    r = kern.mul_mat("q4_K", ws[0], k, k, x)
    r = kern.mul_mat("q4_K", ws[1], k, k, r)
    a2 = kern.mul_mat("q4_K", ws[2], k, k, r)
    c1 = kern.mul_mat("q4_K", ws[4], k, k, x)
    b2 = kern.mul_mat("q4_K", ws[0], k, k, c1)
    assert np.array_equal(u32(outs[0][0].reshape(-1)), u32(a2.cpu().numpy().reshape(-1)))
    assert np.array_equal(u32(outs[0][2].reshape(-1)), u32(c1.cpu().numpy().reshape(-1)))
    assert np.array_equal(u32(outs[0][1].reshape(-1)), u32(b2.cpu().numpy().reshape(-1)))


@pytest.mark.parametrize("t", FUSED_TYPES)
@pytest.mark.parametrize("m,k", FUSED_SHAPES)
def test_mul_mat_persistent_single_node(kern, oracle, monkeypatch, t, m, k):
    """one node through the persistent kernel (option batch_single = 1): bit-identical to the per-node kernel."""
    import b200
    import torch
    w_np = random_blocks(t, m, k, seed=3 * m + k)
    x_np = np.random.default_rng(m + k).uniform(-1, 1, size=(1, k)).astype(np.float32)
    w, x = dev_blocks(kern, w_np), kern.to_device(x_np)
    y = torch.zeros((1, m), dtype=torch.float32, device=kern.device)
    nodes = (b200.MulMatNode * 1)(b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), 1, y.data_ptr()))
    with kern.option("batch_single", 1):
        kern.mul_mat_batch(nodes)
    single = kern.mul_mat(t, w, m, k, x).cpu().numpy().reshape(-1)
    assert np.array_equal(u32(single), u32(y.cpu().numpy().reshape(-1)))
    assert oracle.nmse(oracle.mul_mat(t, w_np, m, k, x_np, mode=1), y.cpu().numpy()) <= NMSE_ORACLE


# ---------------------------------------------------------------------------- full-size (BASELINE.json configs)
FULL = [("q4_0", 4096, 4096), ("q4_K", 4096, 4096), ("q4_K", 11008, 4096), ("q4_K", 4096, 11008),
        ("q5_K", 4096, 11008), ("q6_K", 4096, 11008), ("q6_K", 32000, 4096), ("q8_0", 11008, 4096)]


@pytest.mark.parametrize("t,m,k", FULL)
def test_mul_mat_vec_full_size_sampled_rows(kern, oracle, t, m, k):
    """Full Llama shapes: every output row is checked for finiteness and a seeded sample of 96 rows
    (plus first/last) against the oracle; linearity f(x1 + x2) ~ f(x1) + f(x2) is not exact under q8_1,
    so the size-independent property used is row independence: the result of row r must not depend on
    which other rows are in the matrix (sub-matrix of the sampled rows gives bit-identical values)."""
    w = random_blocks(t, m, k, seed=m + k)
    x = np.random.default_rng(1).uniform(-1, 1, size=(1, k)).astype(np.float32)
    xd = kern.to_device(x)
    scratch = kern.quantize_acts(xd)
    got = kern.mul_mat_vec_q(t, dev_blocks(kern, w), m, k, scratch, 1).cpu().numpy()
    assert np.isfinite(got).all()
    rows = np.unique(np.concatenate([[0, m - 1], np.random.default_rng(2).integers(0, m, size=96)]))
    sub = np.ascontiguousarray(w[rows])
    want = oracle.mul_mat(t, sub, len(rows), k, x, mode=1)
    assert oracle.nmse(want, got[:, rows]) <= NMSE_ORACLE
    got_sub = kern.mul_mat_vec_q(t, dev_blocks(kern, sub), len(rows), k, scratch, 1).cpu().numpy()
    assert np.array_equal(u32(got_sub), u32(got[:, rows]))


# ---------------------------------------------------------------------------- MUL_MAT, batched (tensor cores)
# The batched kernel multiplies bf16-dequantized weights with bf16-rounded activations and accumulates
# in f32 (mmq_tc.cuh).  Bar: test-backend-ops' NMSE gate (5e-4) against exact-activation truth (oracle
# mode 0) — in practice ~5e-6 — and against ggml-cpu's own output where a golden vector exists.
NMSE_TC = 5e-5


@pytest.mark.parametrize("t", QUANT_TYPES)
def test_mul_mat_q_golden(kern, oracle, golden, t):
    w = golden[f"mm_{t}_w"]
    x = golden["mm_x"]
    want = golden[f"mm_{t}_y_n16"]
    got = kern.mul_mat_q(t, dev_blocks(kern, w), 32, 512, kern.to_device(x)).cpu().numpy()
    assert oracle.nmse(want, got) <= NMSE_REF, oracle.nmse(want, got)
    assert oracle.nmse(oracle.mul_mat(t, w, 32, 512, x, mode=0), got) <= NMSE_TC


MMQ_SHAPES = [(128, 256, 16), (16, 256, 9), (130, 512, 33), (300, 1024, 64), (257, 2048, 100), (128, 4096, 128),
              (384, 1024, 129), (200, 768, 257), (64, 11008, 48), (1, 256, 17), (129, 256, 300)]


@pytest.mark.parametrize("t", QUANT_TYPES)
@pytest.mark.parametrize("m,k,n", MMQ_SHAPES)
def test_mul_mat_q_vs_oracle(kern, oracle, t, m, k, n):
    w = random_blocks(t, m, k, seed=300 + m + k)
    x = np.random.default_rng(17 + n).uniform(-1, 1, size=(n, k)).astype(np.float32)
    want = oracle.mul_mat(t, w, m, k, x, mode=0)
    got = kern.mul_mat_q(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy()
    assert got.shape == (n, m)
    assert np.isfinite(got).all()
    assert oracle.nmse(want, got) <= NMSE_TC, oracle.nmse(want, got)


@pytest.mark.parametrize("t", ["q4_0", "q8_0"])
@pytest.mark.parametrize("k", [32, 96, 32 * 37])
def test_mul_mat_q_k_tails(kern, oracle, t, k):
    """k that is not a multiple of 64 / 256: partial slabs and zero-filled tail stages"""
    m, n = 70, 20
    w = random_blocks(t, m, k, seed=k)
    x = np.random.default_rng(k).uniform(-1, 1, size=(n, k)).astype(np.float32)
    got = kern.mul_mat_q(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=0), got) <= NMSE_TC


@pytest.mark.parametrize("t", QUANT_TYPES)
def test_mul_mat_q_unaligned_rows(kern, oracle, monkeypatch, t):
    """weights starting at a 2-byte aligned address with a padded row stride: the re-align path"""
    m, k, n = 140, 1024, 40
    rb = k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]
    w = random_blocks(t, m, k, seed=91)
    lead, stride = 6, rb + 10
    buf = np.zeros(lead + m * stride + 64, dtype=np.uint8)
    for r in range(m):
        buf[lead + r * stride: lead + r * stride + rb] = w[r]
    x = np.random.default_rng(92).uniform(-1, 1, size=(n, k)).astype(np.float32)
    wd = kern.to_device(buf)
    got = kern.mul_mat_q(t, wd[lead:], m, k, kern.to_device(x), w_row_stride=stride).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=0), got) <= NMSE_TC


@pytest.mark.parametrize("bn", [32, 64, 128, 256])
def test_mul_mat_q_every_tile_width(kern, oracle, monkeypatch, bn):
    t, m, k, n = "q4_K", 256, 1024, 300
    w = random_blocks(t, m, k, seed=bn)
    x = np.random.default_rng(bn).standard_normal(size=(n, k)).astype(np.float32)
    with kern.option("mmq_bn", bn):
        got = kern.mul_mat_q(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=0), got) <= NMSE_TC


@pytest.mark.parametrize("t,m,k", [("q8_0", 11008, 4096), ("q4_0", 4096, 11008), ("q4_K", 4096, 4096)])
def test_mul_mat_q_full_size_sampled(kern, oracle, t, m, k):
    """BASELINE config 2 shapes (n = 512): finiteness everywhere, a seeded sample of rows x all columns vs the oracle"""
    n = 512
    w = random_blocks(t, m, k, seed=m + k + 1)
    x = np.random.default_rng(3).uniform(-1, 1, size=(n, k)).astype(np.float32)
    got = kern.mul_mat_q(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy()
    assert np.isfinite(got).all()
    rows = np.unique(np.concatenate([[0, 127, 128, m - 1], np.random.default_rng(4).integers(0, m, size=28)]))
    want = oracle.mul_mat(t, np.ascontiguousarray(w[rows]), len(rows), k, x, mode=0)
    assert oracle.nmse(want, got[:, rows]) <= NMSE_TC


@pytest.mark.parametrize("t", QUANT_TYPES)
@pytest.mark.parametrize("scale", [1e5, 1e-6, 3e30, 1e-30])
def test_mul_mat_q_scale_invariance(kern, oracle, t, scale):
    """the batched path must be as indifferent to magnitude as the reference's per-block q8_1 route
    (src/ggml-cuda/quantize.cu:50-143) and ggml-cpu: activations scaled by 1e5 / 1e-6 / 3e30 / 1e-30, one all-zero
    column, one column with a single outlier.  NMSE per column against exact-activation truth."""
    m, k, n = 200, 2048, 24
    w = random_blocks(t, m, k, seed=77)
    x = np.random.default_rng(78).standard_normal(size=(n, k)).astype(np.float32) * np.float32(scale)
    x[5] = 0.0
    x[6, 100] *= 1000.0
    want = oracle.mul_mat(t, w, m, k, x, mode=0)
    got = kern.mul_mat_q(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy()
    assert np.isfinite(got).all()
    assert not got[5].any()
    for c in range(n):
        if c != 5:
            assert nmse64(want[c], got[c]) <= NMSE_REF, (c, nmse64(want[c], got[c]))
    assert nmse64(want, got) <= NMSE_TC


@pytest.mark.parametrize("t", QUANT_TYPES)
def test_mul_mat_q_adversarial_block_scales(kern, oracle, t):
    """edge scale fields in the weights (0, subnormal d, +-65504, sc = m = 63, Q6_K scale -128) on the batched path:
    finite wherever the truth is finite, NMSE per ROW (a row of huge blocks must not hide a row of tiny ones) and
    overall within the harness gate."""
    m, k, n = 96, 2304, 20
    w = adversarial_blocks(t, m, k, seed=31)
    x = np.random.default_rng(32).standard_normal(size=(n, k)).astype(np.float32)
    want = oracle.mul_mat(t, w, m, k, x, mode=0)
    got = kern.mul_mat_q(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy()
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin)
    for r in range(m):
        f = fin[:, r]
        if f.any() and np.abs(want[f, r]).max() > 0:
            assert oracle.nmse(want[f, r], got[f, r]) <= NMSE_REF, (r, oracle.nmse(want[f, r], got[f, r]))
    assert oracle.nmse(want[fin], got[fin]) <= NMSE_REF


def test_mul_mat_q_matches_mat_vec_route_at_the_switch(kern, oracle):
    """n = 8 takes the q8_1 mat-vec route, n = 9 the tensor-core route: the first 8 columns of both must agree within
    the harness gate even for large activations (no cliff at the switch)."""
    t, m, k = "q4_K", 256, 4096
    w = random_blocks(t, m, k, seed=5)
    x = (np.random.default_rng(6).standard_normal(size=(9, k)) * 3e4).astype(np.float32)
    wd = dev_blocks(kern, w)
    y8 = kern.mul_mat(t, wd, m, k, kern.to_device(x[:8])).cpu().numpy().reshape(8, m)
    y9 = kern.mul_mat(t, wd, m, k, kern.to_device(x)).cpu().numpy().reshape(9, m)
    assert np.isfinite(y9).all()
    assert oracle.nmse(y8, y9[:8]) <= NMSE_REF


# ---------------------------------------------------------------------------- BASELINE configs[3] / configs[4] at full size
@pytest.mark.parametrize("t,m,k", [("q4_K", 28672, 8192), ("q4_K", 8192, 28672), ("q6_K", 4096, 14336), ("q4_K", 14336, 4096)])
def test_configs34_decode_full_size_sampled_rows(kern, oracle, t, m, k):
    """the Llama-3-70B FFN shapes of configs[4] and the Llama-3-8B shapes of configs[3], decode (n = 1 and 8) through the
    list entry point: every output finite, a seeded sample of rows against the oracle, and row independence (the
    row-split property): the sampled rows as a sub-matrix give the same bits."""
    import b200
    import torch
    rng = np.random.default_rng(m + k)
    rows = np.unique(np.concatenate([[0, 127, 128, m - 1], rng.integers(0, m, size=60)]))
    sub = random_blocks(t, len(rows), k, seed=k)
    rb = sub.shape[1]
    wd = torch.randint(0, 256, (m * rb,), dtype=torch.uint8, device=kern.device)      # filler rows: any bytes are legal quants ..
    wv = wd.view(m, rb)
    for off in ((0, 2) if t != "q6_K" else (208,)):                                  # .. with sane fp16 scale fields
        wv[:, off::(144 if t != "q6_K" else 210)] = 0
        wv[:, off + 1::(144 if t != "q6_K" else 210)] = 0x20 if t != "q6_K" else 0x10
    wv[torch.from_numpy(rows).to(kern.device).long()] = kern.to_device(sub)
    for n in (1, 8):
        x_np = rng.uniform(-1, 1, size=(n, k)).astype(np.float32)
        x = kern.to_device(x_np)
        y = torch.zeros((n, m), dtype=torch.float32, device=kern.device)
        kern.mul_mat_batch((b200.MulMatNode * 1)(b200.mul_mat_node(t, wd.data_ptr(), m, k, x.data_ptr(), n, y.data_ptr())))
        got = y.cpu().numpy()
        assert np.isfinite(got).all()
        # n = 1: q8_1 activations (oracle mode 1); n = 8: the MMA list kernel, activations through Q8_K like ggml-cpu (mode 2)
        want = oracle.mul_mat(t, sub, len(rows), k, x_np, mode=1 if n == 1 else 2)
        assert oracle.nmse(want, got[:, rows]) <= NMSE_ORACLE, (n, oracle.nmse(want, got[:, rows]))
        ys = torch.zeros((n, len(rows)), dtype=torch.float32, device=kern.device)
        sd = kern.to_device(sub.reshape(-1))
        kern.mul_mat_batch((b200.MulMatNode * 1)(b200.mul_mat_node(t, sd.data_ptr(), len(rows), k, x.data_ptr(), n, ys.data_ptr())))
        assert np.array_equal(u32(ys.cpu().numpy()), u32(got[:, rows])), n


@pytest.mark.parametrize("t,m,k", [("q4_K", 14336, 4096), ("q6_K", 4096, 14336), ("q4_K", 28672, 8192)])
def test_configs34_prefill_n2048_sampled(kern, oracle, t, m, k):
    """prefill n = 2048 (configs[3]; and the configs[4] gate/up shape): finite everywhere, sampled rows x sampled columns
    against exact-activation truth."""
    import torch
    n = 2048
    rng = np.random.default_rng(m + k + 1)
    rows = np.unique(np.concatenate([[0, 127, 128, m - 1], rng.integers(0, m, size=20)]))
    cols = np.unique(np.concatenate([[0, 255, 256, n - 1], rng.integers(0, n, size=60)]))
    sub = random_blocks(t, len(rows), k, seed=k + 1)
    rb = sub.shape[1]
    wd = torch.zeros((m, rb), dtype=torch.uint8, device=kern.device)
    wd[torch.from_numpy(rows).to(kern.device).long()] = kern.to_device(sub)
    x_np = rng.uniform(-1, 1, size=(n, k)).astype(np.float32)
    got = kern.mul_mat_q(t, wd.view(-1), m, k, kern.to_device(x_np)).cpu().numpy()
    assert np.isfinite(got).all()
    want = oracle.mul_mat(t, sub, len(rows), k, np.ascontiguousarray(x_np[cols]), mode=0)
    assert oracle.nmse(want, got[np.ix_(cols, rows)]) <= NMSE_TC
    other = np.setdiff1d(np.arange(m), rows)[:500]
    assert not got[:, other].any()                      # all-zero weight rows give exact zeros


def test_mul_mat_batch_peers_group_of_one(kern, oracle):
    """b200_mul_mat_batch_peers with a single rank (no peers): the symmetric-region bookkeeping (control block, epochs
    that advance from call to call, the closing barrier) on one GPU; same bits as b200_mul_mat_batch.  The multi-rank
    scatter itself needs >= 2 GPUs (tests/test_gpu_peers.py)."""
    import b200
    import torch
    hidden, ffn = 1024, 3072
    ctrl = int(kern.lib.b200_peer_ctrl_bytes())
    region = torch.zeros(ctrl + 4 * (hidden + 2 * ffn + hidden) + 1024, dtype=torch.uint8, device=kern.device)
    base = region.data_ptr()
    assert base % 128 == 0
    data = region[ctrl:]
    h0 = np.random.default_rng(1).uniform(-1, 1, size=(1, hidden)).astype(np.float32)
    data[:4 * hidden].view(torch.float32).copy_(torch.from_numpy(h0.reshape(-1)))
    offs = {"h0": 0, "g": 4 * hidden, "u": 4 * (hidden + ffn), "d": 4 * (hidden + 2 * ffn)}
    specs = [(ffn, hidden, "h0", "g"), (ffn, hidden, "h0", "u"), (hidden, ffn, "u", "d")]
    ws_np = [random_blocks("q4_K", m, k, seed=60 + i) for i, (m, k, _, _) in enumerate(specs)]
    wd = [dev_blocks(kern, w) for w in ws_np]
    nodes = (b200.MulMatNode * 3)(*[b200.mul_mat_node("q4_K", w.data_ptr(), m, k, base + ctrl + offs[src], 1, base + ctrl + offs[dst])
                                    for (m, k, src, dst), w in zip(specs, wd)])
    pg = b200.peer_group(0, [base], region.numel())
    outs = []
    for rep in range(3):
        data[4 * hidden:].zero_()
        kern.mul_mat_batch_peers(nodes, pg)
        torch.cuda.synchronize()
        outs.append(data[4 * hidden:4 * (2 * hidden + 2 * ffn)].view(torch.float32).cpu().numpy().copy())
    data[4 * hidden:].zero_()
    kern.mul_mat_batch(nodes)
    torch.cuda.synchronize()
    plain = data[4 * hidden:4 * (2 * hidden + 2 * ffn)].view(torch.float32).cpu().numpy().copy()
    for o in outs:
        assert np.array_equal(u32(o), u32(plain))
    assert np.isfinite(plain).all() and np.abs(plain).max() > 0


def test_tensor_peak_micro_kernel_arithmetic(kern):
    """the tcgen05 micro-kernel behind the int8 / bf16 tensor peaks computes what its descriptor says:
    kind::i8 (s8 x s8 -> s32) exactly, kind::f16 with bf16 operands -> f32 exactly on small integers."""
    import torch
    rng = np.random.default_rng(9)
    a = rng.integers(-128, 128, size=(128, 128)).astype(np.int8)
    b = rng.integers(-128, 128, size=(256, 128)).astype(np.int8)
    got = kern.tensor_peak_verify("i8", kern.to_device(a.view(np.uint8)), kern.to_device(b.view(np.uint8))).cpu().numpy()
    want = a.astype(np.int32) @ b.astype(np.int32).T
    assert np.array_equal(got, want)
    af = rng.integers(-8, 9, size=(128, 64)).astype(np.float32)
    bf = rng.integers(-8, 9, size=(256, 64)).astype(np.float32)
    ab = torch.from_numpy(af).to(torch.bfloat16).view(torch.uint8).reshape(128, 128).to(kern.device)
    bb = torch.from_numpy(bf).to(torch.bfloat16).view(torch.uint8).reshape(256, 128).to(kern.device)
    gotf = kern.tensor_peak_verify("bf16", ab, bb).cpu().numpy()
    assert np.array_equal(gotf, af @ bf.T)
    # and the timed mode returns something sane: between 0.2 and 20 peta-ops for int8
    ops = kern.tensor_peak("i8", iters=512, reps=2)
    assert 2e14 < ops < 2e16, ops


# ---------------------------------------------------------------------------- MUL_MAT, batched, the q8_1 int8 tensor-core route
I8_TYPES = ["q4_0", "q8_0", "q4_K", "q5_K"]
I8_SHAPES = [(128, 256, 16), (130, 512, 33), (300, 1024, 64), (257, 2048, 100), (128, 4096, 128), (384, 1024, 129),
             (200, 768, 257), (64, 11008, 48), (1, 256, 17), (129, 256, 300)]


@pytest.mark.parametrize("t", I8_TYPES)
@pytest.mark.parametrize("m,k,n", I8_SHAPES)
def test_mul_mat_q_int8_route_vs_oracle(kern, oracle, t, m, k, n):
    """option mmq_route = 2: activations quantized to q8_1, one tcgen05 kind::i8 MMA per 32-block, block scales applied to
    the drained int32 tiles (mmq_i8.cuh) -- the reference's own arithmetic (src/ggml-cuda/mmq.cuh:796-1092), so the
    result is within f32 summation-order noise of the oracle's q8_1 mode and within the harness gate of exact truth."""
    w = random_blocks(t, m, k, seed=300 + m + k)
    x = np.random.default_rng(17 + n).uniform(-1, 1, size=(n, k)).astype(np.float32)
    with kern.option("mmq_route", 2):
        c0 = kern.launch_count()
        got = kern.mul_mat_q(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy()
        assert kern.launch_count() - c0 == 2          # quantize_mmq_q8_1 + the tile kernel
    assert got.shape == (n, m) and np.isfinite(got).all()
    assert oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=1), got) <= NMSE_ORACLE, oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=1), got)
    assert oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=0), got) <= NMSE_REF


@pytest.mark.parametrize("t", ["q4_0", "q8_0"])
@pytest.mark.parametrize("k,int8", [(32, False), (96, False), (32 * 37, False), (256 * 3, True), (256 * 5 + 0, True)])
def test_mul_mat_q_int8_route_applicability(kern, oracle, t, k, int8):
    """the int8 route needs 16-byte-multiple rows (2-D TMA tiles of raw slabs): k % 256 == 0 for the 18 / 34-byte blocks.
    Other k (tails) silently take the bf16 kernel -- still inside the harness gate, one launch less or more."""
    m, n = 70, 20
    w = random_blocks(t, m, k, seed=k)
    x = np.random.default_rng(k).uniform(-1, 1, size=(n, k)).astype(np.float32)
    with kern.option("mmq_route", 2):
        got = kern.mul_mat_q(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=0), got) <= NMSE_REF
    q8 = oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=1), got)
    assert (q8 <= NMSE_ORACLE) == int8, (k, q8)


@pytest.mark.parametrize("t", I8_TYPES)
def test_mul_mat_q_int8_route_scales_and_adversarial(kern, oracle, t):
    """per-block q8_1 scaling makes the int8 route indifferent to magnitude by construction; edge block scales included"""
    m, k, n = 96, 2304, 40
    w = adversarial_blocks(t, m, k, seed=31)
    x = (np.random.default_rng(32).standard_normal(size=(n, k)) * 1e5).astype(np.float32)
    x[3] = 0.0
    with kern.option("mmq_route", 2):
        got = kern.mul_mat_q(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy()
    want = oracle.mul_mat(t, w, m, k, x, mode=1)
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin)
    assert nmse64(want[fin], got[fin]) <= 1e-9
    assert not got[3].any()


@pytest.mark.parametrize("bn", [32, 64, 128])
def test_mul_mat_q_int8_route_every_tile_width(kern, oracle, bn):
    t, m, k, n = "q4_K", 256, 1024, 300
    w = random_blocks(t, m, k, seed=bn)
    x = np.random.default_rng(bn).standard_normal(size=(n, k)).astype(np.float32)
    with kern.option("mmq_route", 2), kern.option("mmq_bn", bn):
        got = kern.mul_mat_q(t, dev_blocks(kern, w), m, k, kern.to_device(x)).cpu().numpy()
    assert oracle.nmse(oracle.mul_mat(t, w, m, k, x, mode=1), got) <= NMSE_ORACLE
