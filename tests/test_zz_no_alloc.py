"""b200_mul_mat_vec_static: the decode MUL_MAT entry point that allocates nothing (its activation scratch is a static
__device__ array inside the library, csrc/no_alloc.cu).  Called through the C ABI and checked against the existing
pair b200_quantize_acts + b200_mul_mat_vec_q (caller scratch), which runs the same kernels: the results must be
bit-identical (and with that inside the NMSE 1e-9 bar the parity tests use for same-arithmetic comparisons).

Collected last on purpose (file name).
"""
import numpy as np
import pytest

from oracle_lib import QUANT_TYPES, random_blocks

pytestmark = pytest.mark.gpu

NMSE_SAME_ARITHMETIC = 1e-9
ERR_TOO_LARGE = 7


def _nmse64(want, got):
    w, g = np.asarray(want, dtype=np.float64).reshape(-1), np.asarray(got, dtype=np.float64).reshape(-1)
    return float(((w - g) ** 2).sum() / max((w ** 2).sum(), 1e-300))


def _case(kern, t, m, k, n, seed):
    torch = kern.torch
    w = kern.to_device(np.ascontiguousarray(random_blocks(t, m, k, seed)).reshape(-1))
    rng = np.random.default_rng(seed + 7)
    x = kern.to_device(rng.standard_normal((n, k)).astype(np.float32))
    want = kern.mul_mat_vec_q(t, w, m, k, kern.quantize_acts(x), n)
    out = torch.full((n, m), float("nan"), dtype=torch.float32, device=kern.device)
    torch.cuda.synchronize()
    return w, x, want, out


@pytest.mark.parametrize("t", QUANT_TYPES)
def test_static_scratch_matches_caller_scratch_twice_in_a_row(kern, t):
    torch = kern.torch
    # the second call is smaller in both n and k than the first: whatever the first left in the array must not leak in
    cases = [_case(kern, t, 2048, 4096, 8, 101), _case(kern, t, 1536, 2048, 3, 202)]
    # first use on this device: the library looks up the array's address (and CUDA may lazily load the module that holds
    # it), so measure memory around the calls that follow
    w0, x0, _, out0 = _case(kern, t, 64, 256, 1, 11)
    assert kern.mul_mat_vec_static(t, w0, 64, 256, x0, out0) == 0
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    torch_before = torch.cuda.memory_allocated()
    rcs = [kern.mul_mat_vec_static(t, w, want.shape[1], x.shape[1], x, out) for (w, x, want, out) in cases]
    torch.cuda.synchronize()
    assert rcs == [0, 0]
    assert torch.cuda.memory_allocated() == torch_before
    assert torch.cuda.mem_get_info()[0] >= free_before - (2 << 20)       # no device allocation behind the call
    for (_, _, want, out) in cases:
        got, ref = out.cpu().numpy(), want.cpu().numpy()
        assert np.isfinite(got).all()
        assert _nmse64(ref, got) <= NMSE_SAME_ARITHMETIC
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_static_scratch_single_column_and_repeat(kern):
    """n = 1 (the decode shape) and the same call issued twice: identical both times."""
    w, x, want, out = _case(kern, "q4_K", 4096, 4096, 1, 303)
    ref = want.cpu().numpy()
    for _ in range(2):
        out.fill_(float("nan"))
        assert kern.mul_mat_vec_static("q4_K", w, 4096, 4096, x, out) == 0
        assert np.array_equal(out.cpu().numpy().view(np.uint32), ref.view(np.uint32))


def test_static_scratch_rejects_sizes_beyond_the_array(kern):
    torch = kern.torch
    limit = int(kern.lib.b200_static_scratch_bytes())
    assert limit == 1 << 20
    k, n = 262144, 8                                               # 2 MiB of int8 alone
    assert int(kern.lib.b200_acts_scratch_bytes(k, n)) > limit
    dummy = torch.zeros(64, dtype=torch.float32, device=kern.device)
    # rejected before anything is read or launched, so the (too small) buffers are never touched
    rc = int(kern.lib.b200_mul_mat_vec_static(14, dummy.data_ptr(), k // 256 * 210, 4, k, dummy.data_ptr(), k * 4, n,
                                              dummy.data_ptr(), 16, None))
    assert rc == ERR_TOO_LARGE
    assert b"static scratch" in kern.lib.b200_error_string(rc)
    # the largest n = 8 problem that fits does run
    k_fit = 32 * ((limit - 256) // (8 * 36))
    assert int(kern.lib.b200_acts_scratch_bytes(k_fit, 8)) <= limit
