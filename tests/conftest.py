"""pytest configuration: markers, paths and shared fixtures.

`-m "not gpu"`: oracle vs the reference-generated golden vectors, oracle vs oracle/_ref (when it was
built), host logic, and that the C-ABI library loads and exports every declared symbol.
`-m gpu`: the parity tests proper (CUDA path through the C ABI / the ggml plugin vs the oracle).
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "ggml-gfx906_b200")
for p in (PKG, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    path = os.path.join(ROOT, "tests", "golden", "golden_v1.npz")
    return np.load(path)


@pytest.fixture(scope="session")
def golden2():
    """quantize_row_q8_K_ref vectors (oracle/gen_golden.py v2)"""
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_v2.npz"))


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    return oracle_lib.Oracle()


@pytest.fixture(scope="session")
def ref_host():
    """The unmodified reference (oracle/_ref), if it has been built; else skip."""
    import ggml_host
    d = os.path.join(ROOT, "oracle", "_ref", "bin")
    if not os.path.exists(os.path.join(d, "libggml-base.so")):
        pytest.skip("oracle/_ref not built")
    return ggml_host.GgmlHost(d)


@pytest.fixture(scope="session")
def kern():
    """The product kernel library through its C ABI (needs a GPU to compute)."""
    import b200
    return b200.Kernels()
