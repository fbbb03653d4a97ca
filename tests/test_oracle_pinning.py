"""Pin the CPU oracle against the unmodified reference compiled into oracle/_ref (live calls into
the reference's own dequantize_row_* / quantize_row_q8_1_ref / ggml_quantize_chunk), on much larger
random inputs than the committed fixtures.  Skips where oracle/_ref is absent."""
import numpy as np
import pytest

from oracle_lib import BLOCK_BYTES, BLOCK_ELEMS, EXTRA_TYPES, QUANT_TYPES, random_blocks


@pytest.mark.parametrize("t", QUANT_TYPES + EXTRA_TYPES)
def test_dequantize_random_bytes(oracle, ref_host, t):
    # raw random bytes incl. arbitrary fp16 scale bit patterns (inf/nan compare by bits)
    rng = np.random.default_rng(7)
    k = 256 * 64
    for rep in range(4):
        blocks = rng.integers(0, 256, size=k // BLOCK_ELEMS[t] * BLOCK_BYTES[t], dtype=np.uint8)
        want = ref_host.dequantize(t, blocks, k)
        got = oracle.dequantize(t, blocks, k)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (t, rep)


@pytest.mark.parametrize("t", QUANT_TYPES + EXTRA_TYPES)
def test_dequantize_quantized_weights(oracle, ref_host, t):
    rng = np.random.default_rng(11)
    k = 1024
    x = rng.uniform(-1, 1, size=(32, k)).astype(np.float32)
    q = ref_host.quantize(t, x)
    for r in range(q.shape[0]):
        want = ref_host.dequantize(t, q[r], k)
        got = oracle.dequantize(t, q[r], k)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        assert np.array_equal(oracle.dequantize_f16(t, q[r], k), ref_host.fp32_to_fp16(want))


def test_quantize_q8_1_random(oracle, ref_host):
    rng = np.random.default_rng(13)
    for scale in (1.0, 1e-3, 1e3, 1e-30, 1e30):
        x = (rng.standard_normal(32 * 4096) * scale).astype(np.float32)
        assert np.array_equal(oracle.quantize_q8_1(x), ref_host.quantize_q8_1(x)), scale
    # every f32 bit pattern class, NaN excluded (MAX() drops NaN in amax but roundf(NaN) is UB-cast)
    bits = rng.integers(0, 2**32, size=32 * 8192, dtype=np.uint64).astype(np.uint32).view(np.float32)
    bits = np.where(np.isfinite(bits), bits, np.float32(0.5))
    assert np.array_equal(oracle.quantize_q8_1(bits), ref_host.quantize_q8_1(bits))


@pytest.mark.parametrize("t", ("q4_0", "q8_0"))
def test_weight_quantizers(oracle, ref_host, t):
    rng = np.random.default_rng(17)
    x = rng.uniform(-1, 1, size=(64, 512)).astype(np.float32)
    assert np.array_equal(oracle.quantize(t, x), ref_host.quantize(t, x))


@pytest.mark.parametrize("t", QUANT_TYPES + EXTRA_TYPES)
def test_random_blocks_are_sane(oracle, t):
    b = random_blocks(t, 4, 1024, seed=3)
    y = np.stack([oracle.dequantize(t, b[r], 1024) for r in range(4)])
    assert np.isfinite(y).all() and 0.05 < np.abs(y).max() < 64.0


@pytest.mark.parametrize("t", QUANT_TYPES + EXTRA_TYPES)
def test_mul_mat_live_vs_reference_cpu(oracle, ref_host, t):
    rng = np.random.default_rng(19)
    m, k, n = 48, 1024, 5
    w = ref_host.quantize(t, rng.uniform(-1, 1, size=(m, k)).astype(np.float32))
    x = rng.uniform(-1, 1, size=(n, k)).astype(np.float32)
    cpu = ref_host.cpu_backend(2)
    try:
        want = ref_host.mul_mat(cpu, t, w, m, k, x)
    finally:
        ref_host.base.ggml_backend_free(cpu)
    got = oracle.mul_mat(t, w, m, k, x, mode=1)
    assert oracle.nmse(want, got) < 5e-4


def test_quantize_q8_K_random(oracle, ref_host):
    rng = np.random.default_rng(23)
    for scale in (1.0, 1e-3, 1e3, 1e-30, 1e30):
        x = (rng.standard_normal(256 * 512) * scale).astype(np.float32)
        x[rng.integers(0, x.size, size=64)] = 0.0
        q, d, bs = oracle.quantize_q8_K(x)
        rq, rd, rbs = ref_host.quantize_q8_K(x)
        assert np.array_equal(q, rq) and np.array_equal(d.view(np.uint32), rd.view(np.uint32)) and np.array_equal(bs, rbs), scale
    # magnitude ties with opposite signs, all-equal blocks
    x = np.tile(np.array([1.0, -1.0], dtype=np.float32), 128)
    for v in (x, -x, np.full(256, 3.0, dtype=np.float32), np.full(256, -3.0, dtype=np.float32)):
        q, d, bs = oracle.quantize_q8_K(v)
        rq, rd, rbs = ref_host.quantize_q8_K(v)
        assert np.array_equal(q, rq) and np.array_equal(d.view(np.uint32), rd.view(np.uint32)) and np.array_equal(bs, rbs)


@pytest.mark.parametrize("t", ("q4_K", "q5_K", "q6_K"))
def test_mul_mat_q8_K_mode_live_vs_reference_cpu(oracle, ref_host, t):
    rng = np.random.default_rng(29)
    m, k, n = 48, 2048, 5
    w = ref_host.quantize(t, rng.uniform(-1, 1, size=(m, k)).astype(np.float32))
    x = (rng.standard_normal((n, k)) * 3).astype(np.float32)
    cpu = ref_host.cpu_backend(2)
    try:
        want = ref_host.mul_mat(cpu, t, w, m, k, x)
    finally:
        ref_host.base.ggml_backend_free(cpu)
    assert oracle.nmse(want, oracle.mul_mat(t, w, m, k, x, mode=2)) < 1e-10
