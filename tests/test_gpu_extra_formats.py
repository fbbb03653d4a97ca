"""Q4_1 / Q5_0 / Q5_1 / IQ4_NL ("the remaining quant formats") through the kernel library's general kernels --
dequantize (f32 / f16 / bf16), GET_ROWS, strided CPY and the one-warp-per-row mat-vec behind MUL_MAT / MUL_MAT_ID -- and the bf16
output of the dequantizers for every format (reference: src/ggml-cuda/convert.cu:650-659).  Dequantization bit-exact against
the oracle (pinned to the reference's dequantize_row_* live and by golden_v2.npz), MUL_MAT against the oracle's q8_1 arithmetic."""
import numpy as np
import pytest

from oracle_lib import EXTRA_TYPES, QUANT_TYPES, random_blocks

pytestmark = pytest.mark.gpu

NMSE_ORACLE = 1e-9


def u32(a):
    return np.ascontiguousarray(a).view(np.uint32)


def dev(kern, blocks):
    return kern.to_device(np.ascontiguousarray(blocks).reshape(-1))


@pytest.mark.parametrize("t", EXTRA_TYPES)
def test_dequantize_golden(kern, golden2, t):
    blocks = golden2[f"{t}_blocks"]
    n = blocks.shape[0] * 512
    got32 = kern.dequantize_row(t, dev(kern, blocks), n, "f32").cpu().numpy()
    assert np.array_equal(u32(got32), u32(golden2[f"{t}_deq_f32"].reshape(-1)))
    got16 = kern.dequantize_row(t, dev(kern, blocks), n, "f16").cpu().numpy()
    assert np.array_equal(got16.view(np.uint16), golden2[f"{t}_deq_f16"].reshape(-1))


@pytest.mark.parametrize("t", EXTRA_TYPES)
def test_dequantize_and_get_rows_vs_oracle(kern, oracle, t):
    import torch
    nrows, k = 96, 2048 + 32
    blocks = random_blocks(t, nrows, k, seed=3)
    # (finite fp16 scale fields only: NaN payloads propagate differently through x86 and GPU arithmetic; the golden vectors
    #  above carry the finite edge patterns: zero, -0, subnormals, the largest finite half)
    want = oracle.dequantize(t, blocks, nrows * k)
    got = kern.dequantize_row(t, dev(kern, blocks), nrows * k, "f32").cpu().numpy()
    assert np.array_equal(u32(got), u32(want))
    rows = np.array([5, 0, 95, 5, 17], dtype=np.int32)
    gr = kern.get_rows(t, dev(kern, blocks), k, torch.from_numpy(rows).to(kern.device)).cpu().numpy()
    assert np.array_equal(u32(gr), u32(want.reshape(nrows, k)[rows]))
    # ragged size (not a multiple of the 256-element CTA span)
    got1 = kern.dequantize_row(t, dev(kern, blocks[:1]), k, "f32").cpu().numpy()
    assert np.array_equal(u32(got1), u32(want[:k]))


@pytest.mark.parametrize("t", QUANT_TYPES + EXTRA_TYPES)
def test_dequantize_to_bf16(kern, oracle, t):
    """bf16 output = RN-even narrowing of the reference's f32 values (ggml_fp32_to_bf16_row semantics for finite values)"""
    import torch
    nrows, k = 16, 4096
    blocks = random_blocks(t, nrows, k, seed=8)
    want = torch.from_numpy(oracle.dequantize(t, blocks, nrows * k)).to(torch.bfloat16)
    got = kern.dequantize_row(t, dev(kern, blocks), nrows * k, "bf16").cpu()
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))


@pytest.mark.parametrize("t", EXTRA_TYPES)
@pytest.mark.parametrize("n", [1, 3, 8, 21])
def test_mul_mat_vs_oracle(kern, oracle, t, n):
    """n <= 8: one launch of the general mat-vec; n = 21: groups of 8 columns; also through the list entry point"""
    import b200
    import torch
    m, k = 200, 1024 + 64
    w_np = random_blocks(t, m, k, seed=20)
    x_np = np.random.default_rng(21).standard_normal((n, k)).astype(np.float32)
    w, x = dev(kern, w_np), kern.to_device(x_np)
    want = oracle.mul_mat(t, w_np, m, k, x_np, mode=1)
    got = kern.mul_mat(t, w, m, k, x).cpu().numpy().reshape(n, m)
    assert oracle.nmse(want, got) <= NMSE_ORACLE, (t, n)
    y = torch.zeros((n, m), dtype=torch.float32, device=kern.device)
    kern.mul_mat_batch((b200.MulMatNode * 2)(b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), n, y.data_ptr()),
                                            b200.mul_mat_node("q4_K", dev(kern, random_blocks("q4_K", 8, 1024, 1)).data_ptr(), 8, 1024,
                                                              kern.to_device(x_np[:, :1024].copy()).data_ptr(), n, torch.zeros((n, 8), device=kern.device).data_ptr())))
    assert np.array_equal(u32(y.cpu().numpy()), u32(got))


@pytest.mark.parametrize("t", EXTRA_TYPES)
def test_mul_mat_broadcast_channels(kern, oracle, t):
    """4-D: weights [k, m, 2, 1], activations [k, n, 4, 3] with n > 8 (channel broadcast ratios 2 and 3)"""
    m, k, n = 33, 256, 11
    ne_a, ne_b = [k, m, 2, 1], [k, n, 4, 3]
    rb = k // 32 * len(random_blocks(t, 1, 32, 0)[0])
    w_np = np.stack([random_blocks(t, m, k, seed=30 + c) for c in range(2)])
    x_np = np.random.default_rng(31).uniform(-1, 1, size=(3, 4, n, k)).astype(np.float32)
    nb_a = [rb // (k // 32), rb, rb * m, rb * m * 2]
    nb_b = [4, 4 * k, 4 * k * n, 4 * k * n * 4]
    out = kern.mul_mat_4d(t, dev(kern, w_np), ne_a, nb_a, kern.to_device(x_np), ne_b, nb_b).cpu().numpy()
    for i13 in range(3):
        for i12 in range(4):
            want = oracle.mul_mat(t, w_np[i12 // 2], m, k, x_np[i13, i12], mode=1)
            assert oracle.nmse(want, out[i13, i12]) <= NMSE_ORACLE, (t, i12, i13)


@pytest.mark.parametrize("t", ("q8_0", "q4_0"))
def test_quantize_rows_bit_exact(kern, oracle, golden, t):
    """f32 -> Q8_0 / Q4_0 blocks (the CPY f32 -> quant direction): bit for bit quantize_row_*_ref -- the golden rows the
    reference quantized without an importance matrix, random rows through the oracle, ties / zeros / one outlier, strided input"""
    src = golden[f"{t}_src_f32"][:3]
    got = kern.quantize_rows(t, kern.to_device(np.ascontiguousarray(src))).cpu().numpy()
    assert np.array_equal(got, golden[f"{t}_blocks"][:3])
    rng = np.random.default_rng(5)
    x = rng.standard_normal((37, 1024 + 32)).astype(np.float32) * 3
    x[1] = 0.0
    x[2, :64] = np.tile(np.array([1.0, -1.0], dtype=np.float32), 32)      # |max| ties of opposite sign: the first one sets the scale
    x[3, :32] = (np.arange(32) - 15.5).astype(np.float32)                 # x.5 ties after scaling
    x[4, 7] = 1e6
    big = kern.to_device(np.concatenate([x, x], axis=1))
    got = kern.quantize_rows(t, big[:, :x.shape[1]]).cpu().numpy()        # row stride twice the row
    assert np.array_equal(got, oracle.quantize(t, x))


@pytest.mark.parametrize("t", ("q8_0", "q4_0"))
def test_quantize_rows_4d_permuted_source(kern, oracle, t):
    """CPY f32 -> quant from a permuted source (ggml permute 0,2,1,3: dims 1 and 2 swapped) into a dense destination: every row
    (i1, i2, i3) bit for bit the quantized source row"""
    from b200 import row_bytes
    rng = np.random.default_rng(9)
    k, a, b, c = 256, 2, 3, 4                       # the view has ne = [k, a, b, c]; its memory is [c][a][b][k]
    mem = rng.standard_normal((c, a, b, k)).astype(np.float32)
    view = np.ascontiguousarray(mem.transpose(0, 2, 1, 3))      # [c][b][a][k]: what the view reads in index order
    rb = row_bytes(t, k)
    nb_x = [4, 4 * k * b, 4 * k, 4 * k * b * a]     # strides of ne[1] = a, ne[2] = b, ne[3] = c in mem
    nb_y = [rb // (k // 32), rb, rb * a, rb * a * b]
    got = kern.quantize_rows_4d(t, kern.to_device(mem), [k, a, b, c], nb_x, nb_y, rb * a * b * c).cpu().numpy()
    want = oracle.quantize(t, view.reshape(-1, k))
    assert np.array_equal(got.reshape(-1, rb), want)


@pytest.mark.parametrize("t", tuple(QUANT_TYPES) + ("q5_1",))
@pytest.mark.parametrize("m,k,n_tok", [(300, 1024, 1), (33, 4096, 3), (4100, 512, 2)])
def test_mul_mat_id_vs_oracle_and_bad_expert(kern, oracle, t, m, k, n_tok):
    """MUL_MAT_ID through the C ABI (mixture of experts; ggml_mul_mat_id, src/ggml-cpu/ggml-cpu.c:1500-1620): every (slot, token)
    channel is the one-column mat-vec of the expert its id names -- NMSE 1e-9 against the oracle's q8_1 arithmetic per channel,
    whichever kernel serves it (TMA-staged stream kernel / one warp per row); an id outside the stack zero-fills its channel and
    raises the sticky error flag (the reference asserts)."""
    import b200
    import torch
    n_as, n_used = 4, 2
    rng = np.random.default_rng(31)
    w_np = [random_blocks(t, m, k, seed=60 + e) for e in range(n_as)]
    experts = kern.to_device(np.concatenate([w.reshape(-1) for w in w_np]))
    x_np = rng.uniform(-1, 1, size=(n_tok, 1, k)).astype(np.float32)
    ids_np = np.stack([rng.permutation(n_as)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    assert kern.device_error_flags() == 0
    got = kern.mul_mat_id(t, experts, m, k, n_as, kern.to_device(x_np), torch.from_numpy(ids_np).to(kern.device)).cpu().numpy()
    assert kern.device_error_flags() == 0
    for tok in range(n_tok):
        for e in range(n_used):
            want = oracle.mul_mat(t, w_np[ids_np[tok, e]], m, k, x_np[tok], mode=1)
            assert oracle.nmse(want, got[tok, e].reshape(1, m)) <= NMSE_ORACLE, (t, tok, e)
    bad = ids_np.copy()
    bad[0, 1] = n_as
    got2 = kern.mul_mat_id(t, experts, m, k, n_as, kern.to_device(x_np), torch.from_numpy(bad).to(kern.device)).cpu().numpy()
    assert kern.device_error_flags() & b200.ERR_FLAG_INDEX
    assert not got2[0, 1].any()
    assert np.array_equal(got2[0, 0], got[0, 0])
