"""CPU-side checks of the drop-in boundary: the C-ABI libraries load and export every symbol that
include/*.h declares; argument validation that needs no device works; the product tree never
references the oracle."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "ggml-gfx906_b200")
KLIB = os.path.join(PKG, "libggml_b200_kernels.so")
PLUGIN = os.path.join(PKG, "libggml-b200.so")


def declared_symbols(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"^\s*#.*$", "", src, flags=re.M)
    return sorted(set(re.findall(r"B200_API\s+[^;(]*?\b(\w+)\s*\(", src)))


@pytest.fixture(scope="module")
def built():
    if not os.path.exists(KLIB):
        import __graft_entry__ as g
        g.build()
    return True


def test_kernel_library_exports_every_declared_symbol(built):
    lib = C.CDLL(KLIB)
    syms = declared_symbols("ggml_b200.h")
    assert len(syms) >= 18, syms
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/ggml_b200.h but not exported"


def test_plugin_exports_every_declared_symbol(built):
    if not os.path.exists(PLUGIN):
        pytest.skip("plugin not built (needs the reference headers at build time)")
    base = os.path.join(ROOT, "oracle", "_ref", "bin", "libggml-base.so")
    if not os.path.exists(base):
        pytest.skip("the plugin resolves ggml-base symbols from its host; oracle/_ref not built")
    C.CDLL(base, mode=C.RTLD_GLOBAL)      # the host application's ggml, as in a real load
    lib = C.CDLL(PLUGIN)
    syms = declared_symbols("ggml_b200_plugin.h")
    assert "ggml_backend_init" in syms and "ggml_backend_score" in syms
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/ggml_b200_plugin.h but not exported"


def test_format_table_and_errors(built):
    import b200
    lib = b200.load_library()
    assert lib.b200_abi_version() == 2
    for t in b200.QUANT_TYPES:
        assert lib.b200_block_elems(b200.TYPE_ID[t]) == b200.BLOCK_ELEMS[t]
        assert lib.b200_block_bytes(b200.TYPE_ID[t]) == b200.BLOCK_BYTES[t]
        assert lib.b200_row_bytes(b200.TYPE_ID[t], 4096) == b200.row_bytes(t, 4096)
        assert lib.b200_row_bytes(b200.TYPE_ID[t], 4097) == 0
    # argument validation happens before any CUDA call
    assert lib.b200_dequantize_row(b200.TYPE_ID["f32"], None, None, 32, 0, None) == 1      # B200_ERR_TYPE
    assert lib.b200_dequantize_row(b200.TYPE_ID["q4_K"], None, None, 100, 0, None) == 2    # B200_ERR_SHAPE
    assert lib.b200_mul_mat_vec_q(b200.TYPE_ID["q4_K"], None, 144, 4, 256, None, 9, None, 16, None) == 2
    assert lib.b200_error_string(2) == b"bad shape"
    assert lib.b200_acts_scratch_bytes(4096, 1) == 4096 + 128 * 4


def test_batch_workspace_size_covers_every_list_path(built):
    """b200_mul_mat_batch_workspace_size is host arithmetic: it must cover what the list paths carve out of the
    workspace -- one q8 record (k + k/8 bytes) per one-column node for the persistent kernel, the three activation
    planes of the MMA list kernel (8 k int8 + k/32 x 8 x (8 + 8) bytes) per node with 2..8 columns, and the
    largest single-node need on top -- and grow with the list."""
    import b200
    lib = b200.load_library()

    def size(nodes):
        arr = (b200.MulMatNode * len(nodes))(*nodes)
        return int(lib.b200_mul_mat_batch_workspace_size(arr, len(nodes)))

    def node(t, m, k, n):
        return b200.mul_mat_node(t, 0x10000, m, k, 0x20000, n, 0x30000)

    assert size([]) >= 256
    for k in (4096, 14336):
        one = size([node("q4_K", 4096, k, 1)])
        assert one >= k + k // 8
        assert size([node("q4_K", 4096, k, 1)] * 10) >= 10 * (k + k // 8)
        for n in (2, 4, 8):
            for t in ("q4_K", "q5_K", "q6_K"):
                assert size([node(t, 4096, k, n)]) >= 8 * k + (k // 256) * 8 * 52, (t, k, n)      # int8 columns + Q8_K scales and sums
            assert size([node("q4_K", 4096, k, n), node("q6_K", 1024, k, n)]) >= 2 * (8 * k + (k // 256) * 8 * 52)
            assert size([node("q8_0", 4096, k, n)]) >= n * (k + k // 8)          # one persistent entry per column
        big = size([node("q4_K", 4096, k, 512)])
        assert big >= 512 * k                                                     # batched path: activations staged once
        assert size([node("q4_K", 4096, k, 512), node("q4_K", 4096, k, 1)]) >= big + k + k // 8


def test_no_device_fails_loudly(built):
    import torch
    import b200
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(b200.B200Error):
        b200.Kernels()


def test_product_never_touches_the_oracle():
    bad = []
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"b200_oracle|oracle_lib|b200o_", txt):
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad
