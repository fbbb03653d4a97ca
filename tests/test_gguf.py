"""GGUF -> device: the reader half of csrc/gguf_load.cu on the CPU, pinned to the reference's own
gguf_init_from_file (oracle/_ref: src/gguf.cpp:319-735) on the same files — accepted files give the same keys, values,
tensor infos and offsets; the broken ones are refused by both — and the loader half on the GPU (bytes on the device ==
bytes in the file; a loaded tensor is usable by MUL_MAT)."""
import ctypes as C
import os
import struct

import numpy as np
import pytest

import b200
import gguf_files as gf

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class _GgufInitParams(C.Structure):
    _fields_ = [("no_alloc", C.c_bool), ("ctx", C.c_void_p)]


@pytest.fixture(scope="module")
def ref_gguf():
    """the reference's gguf API from oracle/_ref/bin/libggml-base.so, or None where it is not built"""
    p = os.path.join(ROOT, "oracle", "_ref", "bin", "libggml-base.so")
    if not os.path.exists(p):
        return None
    lib = C.CDLL(p, mode=C.RTLD_GLOBAL)
    vp, i64 = C.c_void_p, C.c_int64
    lib.gguf_init_from_file.restype = vp;  lib.gguf_init_from_file.argtypes = [C.c_char_p, _GgufInitParams]
    lib.gguf_free.restype = None;          lib.gguf_free.argtypes = [vp]
    for name, rt in (("gguf_get_version", C.c_uint32), ("gguf_get_alignment", C.c_size_t), ("gguf_get_data_offset", C.c_size_t),
                     ("gguf_get_n_kv", i64), ("gguf_get_n_tensors", i64)):
        f = getattr(lib, name); f.restype = rt; f.argtypes = [vp]
    for name, rt in (("gguf_get_key", C.c_char_p), ("gguf_get_kv_type", C.c_int), ("gguf_get_arr_type", C.c_int),
                     ("gguf_get_arr_n", C.c_size_t), ("gguf_get_tensor_offset", C.c_size_t), ("gguf_get_tensor_name", C.c_char_p),
                     ("gguf_get_tensor_type", C.c_int), ("gguf_get_tensor_size", C.c_size_t), ("gguf_get_val_str", C.c_char_p),
                     ("gguf_get_val_data", vp), ("gguf_get_arr_data", vp)):
        f = getattr(lib, name); f.restype = rt; f.argtypes = [vp, i64]
    lib.gguf_get_arr_str.restype = C.c_char_p; lib.gguf_get_arr_str.argtypes = [vp, i64, C.c_size_t]
    lib.ggml_type_size.restype = C.c_size_t;   lib.ggml_type_size.argtypes = [C.c_int]
    lib.ggml_blck_size.restype = i64;          lib.ggml_blck_size.argtypes = [C.c_int]
    return lib


def ref_open(lib, path):
    return lib.gguf_init_from_file(os.fsencode(path), _GgufInitParams(True, None))


@pytest.fixture(scope="module")
def klib():
    return b200.load_library()


def test_reads_what_was_written(tmp_path, klib):
    kvs, tensors = gf.sample_kvs(), gf.sample_tensors()
    p = str(tmp_path / "a.gguf")
    blob = gf.write(p, kvs, tensors)
    with b200.Gguf(p, klib) as g:
        assert g.version == 3 and g.alignment == 32
        assert g.keys() == [k[0] for k in kvs]
        for kv in kvs:
            got = g.value(kv[0])
            if kv[1] == gf.ARR:
                assert list(got) == list(kv[2]), kv[0]
            elif kv[1] in (gf.F32,):
                assert got == np.float32(kv[2])
            else:
                assert got == kv[2], kv[0]
        infos = g.tensors()
        assert [t["name"] for t in infos] == [t[0] for t in tensors]
        for info, (name, ne, tid, raw) in zip(infos, tensors):
            assert info["type"] == tid and info["n_dims"] == len(ne) and info["ne"] == list(ne) + [1] * (4 - len(ne))
            assert info["nbytes"] == len(raw)
            assert blob[info["file_offset"]:info["file_offset"] + info["nbytes"]] == raw
            assert info["file_offset"] % 32 == 0 and info["device_offset"] % 256 == 0
        assert g.data_offset + g.data_bytes == len(blob)
        assert g.find_tensor("blk.0.ffn_down.weight") == 3 and g.find_tensor("nope") == -1
        with pytest.raises(KeyError):
            g.value("nope")
        # device image: in file order, 256-byte aligned, no overlap
        end = 0
        for info in infos:
            assert info["device_offset"] >= end
            end = info["device_offset"] + info["nbytes"]
        assert g.device_bytes >= end


@pytest.mark.parametrize("alignment", [32, 64, 4096])
def test_same_view_as_the_reference_reader(tmp_path, klib, ref_gguf, alignment):
    if ref_gguf is None:
        pytest.skip("oracle/_ref not built")
    kvs, tensors = gf.sample_kvs(), gf.sample_tensors(seed=alignment)
    p = str(tmp_path / "a.gguf")
    gf.write(p, kvs, tensors, alignment=alignment)
    r = ref_open(ref_gguf, p)
    assert r, "the reference refuses the fixture writer's file"
    try:
        with b200.Gguf(p, klib) as g:
            assert g.version == ref_gguf.gguf_get_version(r)
            assert g.alignment == ref_gguf.gguf_get_alignment(r) == alignment
            assert g.data_offset == ref_gguf.gguf_get_data_offset(r)
            n_kv = ref_gguf.gguf_get_n_kv(r)
            assert len(g.keys()) == n_kv
            for i in range(n_kv):
                key = ref_gguf.gguf_get_key(r, i).decode()
                assert g.keys()[i] == key
                t = ref_gguf.gguf_get_kv_type(r, i)
                assert klib.b200_gguf_kv_type(g.h, i) == t
                if t == gf.ARR:
                    et, n = ref_gguf.gguf_get_arr_type(r, i), ref_gguf.gguf_get_arr_n(r, i)
                    assert klib.b200_gguf_arr_type(g.h, i) == et and klib.b200_gguf_arr_n(g.h, i) == n
                    if et == gf.STR:
                        assert g.value(key) == [ref_gguf.gguf_get_arr_str(r, i, j).decode() for j in range(n)]
                    elif n:
                        size = n * np.dtype(b200.Gguf._NP[et]).itemsize
                        assert C.string_at(ref_gguf.gguf_get_arr_data(r, i), size) == C.string_at(klib.b200_gguf_val_data(g.h, i), size)
                elif t == gf.STR:
                    assert g.value(key) == ref_gguf.gguf_get_val_str(r, i).decode()
                else:
                    size = np.dtype(b200.Gguf._NP[t]).itemsize
                    assert C.string_at(ref_gguf.gguf_get_val_data(r, i), size) == C.string_at(klib.b200_gguf_val_data(g.h, i), size)
            infos = g.tensors()
            assert len(infos) == ref_gguf.gguf_get_n_tensors(r)
            for i, info in enumerate(infos):
                assert info["name"] == ref_gguf.gguf_get_tensor_name(r, i).decode()
                assert info["type"] == ref_gguf.gguf_get_tensor_type(r, i)
                assert info["nbytes"] == ref_gguf.gguf_get_tensor_size(r, i)
                assert info["file_offset"] == ref_gguf.gguf_get_data_offset(r) + ref_gguf.gguf_get_tensor_offset(r, i)
    finally:
        ref_gguf.gguf_free(r)


def test_type_table_matches_the_reference(klib, ref_gguf):
    if ref_gguf is None:
        pytest.skip("oracle/_ref not built")
    for t in range(40):
        qk, ts = ref_gguf.ggml_blck_size(t), ref_gguf.ggml_type_size(t)
        if qk == 0:
            assert klib.b200_gguf_type_row_bytes(t, 256) == 0, t
        else:
            assert klib.b200_gguf_type_row_bytes(t, 5 * 256) == 5 * 256 // qk * ts, t
            if qk > 1:
                assert klib.b200_gguf_type_row_bytes(t, qk + 1) == 0
    assert klib.b200_gguf_type_row_bytes(40, 256) == 0 and klib.b200_gguf_type_row_bytes(-1, 256) == 0


def _broken_files():
    kvs, tensors = gf.sample_kvs()[:4], gf.sample_tensors()[:3]
    good = gf.build(kvs, tensors)
    yield "bad magic", gf.build(kvs, tensors, magic=b"GGUX")
    yield "empty file", b""
    yield "magic only", b"GGUF"
    yield "version 0", gf.build(kvs, tensors, version=0)
    yield "version 1", gf.build(kvs, tensors, version=1)
    yield "version 4", gf.build(kvs, tensors, version=4)
    yield "byte-swapped version", gf.build(kvs, tensors, version=0x03000000)
    yield "negative tensor count", gf.build(kvs, tensors, n_tensors=-1)
    yield "negative kv count", gf.build(kvs, tensors, n_kv=-5)
    yield "more kvs than the file holds", gf.build(kvs, tensors, n_kv=len(kvs) + 1000)
    yield "more tensors than the file holds", gf.build(kvs, [], n_tensors=3)
    yield "duplicate key", gf.build(kvs + [kvs[0]], tensors)
    yield "invalid value type", _with_raw_kv(kvs, tensors, gf.gstr("bad") + struct.pack("<i", 13) + b"\0" * 8)
    yield "array of arrays", _with_raw_kv(kvs, tensors, gf.gstr("aa") + struct.pack("<iiQ", gf.ARR, gf.ARR, 0))
    yield "alignment 0", gf.build(kvs + [("general.alignment", gf.U32, 0)], tensors)
    yield "alignment 48", gf.build(kvs + [("general.alignment", gf.U32, 48)], tensors)
    yield "duplicate tensor name", gf.build(kvs, tensors + [tensors[0]])
    yield "tensor name of 64 bytes", gf.build(kvs, [("n" * 64, [4], 0, b"\0" * 16)])
    yield "5 dimensions", gf.build(kvs, [("t", [2, 2, 2, 2, 2], 0, b"\0" * 128)])
    yield "negative extent", gf.build(kvs, [("t", [4, -2], 0, b"")])
    yield "element count overflow", gf.build(kvs, [("t", [1 << 40, 1 << 40], 0, b"")])
    yield "removed type id 4", gf.build(kvs, [("t", [32], 4, b"\0" * 32)])
    yield "type id 40", gf.build(kvs, [("t", [32], 40, b"\0" * 32)])
    yield "negative type id", gf.build(kvs, [("t", [32], -1, b"\0" * 32)])
    yield "row not a multiple of the block", gf.build(kvs, [("t", [100, 2], 12, b"\0" * 144)])
    yield "wrong data offset", gf.build(kvs, tensors, offsets=[0, 32, 64])
    yield "string length beyond the file", good[:24] + struct.pack("<Q", 1 << 40) + good[32:]
    yield "cut inside the key/value pairs", good[:60]
    yield "cut inside the tensor infos", good[:good.index(b"blk.0.attn_q.weight") + 8]


def _with_raw_kv(kvs, tensors, raw):
    """a file whose LAST key/value pair is the given raw bytes"""
    b = gf.build(kvs, tensors)
    n_kv_pos = 4 + 4 + 8
    body_kv = b"".join(gf.kv_bytes(*kv) for kv in kvs)
    start = n_kv_pos + 8
    assert b[start:start + len(body_kv)] == body_kv
    return b[:n_kv_pos] + struct.pack("<q", len(kvs) + 1) + body_kv + raw + b[start + len(body_kv):]


@pytest.mark.parametrize("what,blob", list(_broken_files()), ids=[w for w, _ in _broken_files()])
def test_refuses_what_the_reference_refuses(tmp_path, klib, ref_gguf, what, blob):
    p = str(tmp_path / "bad.gguf")
    with open(p, "wb") as f:
        f.write(blob)
    with pytest.raises(b200.B200Error) as ei:
        b200.Gguf(p, klib)
    assert "code 8" in str(ei.value) and len(str(ei.value)) > 30      # B200_ERR_FILE with the reader's reason
    if ref_gguf is not None and what not in ("element count overflow",):
        # (the reference's overflow test divides by ne[1..3]: fine here, but a zero extent there is a SIGFPE we do not provoke)
        r = ref_open(ref_gguf, p)
        if r:
            ref_gguf.gguf_free(r)
        assert not r, f"the reference accepts a file this reader refuses: {what}"


def test_truncated_data_section(tmp_path, klib):
    """header intact, tensor data cut short: refused at open (the reference finds it when it reads the blob, gguf.cpp:700-708)"""
    blob = gf.build(gf.sample_kvs(), gf.sample_tensors())
    p = str(tmp_path / "cut.gguf")
    with open(p, "wb") as f:
        f.write(blob[:-1000])
    with pytest.raises(b200.B200Error):
        b200.Gguf(p, klib)
    with pytest.raises(b200.B200Error):
        b200.Gguf(str(tmp_path / "missing.gguf"), klib)


def test_no_tensors_and_no_kvs(tmp_path, klib, ref_gguf):
    p = str(tmp_path / "empty.gguf")
    gf.write(p, [], [])
    with b200.Gguf(p, klib) as g:
        assert g.keys() == [] and g.tensors() == [] and g.device_bytes == 0
    if ref_gguf is not None:
        r = ref_open(ref_gguf, p)
        assert r
        ref_gguf.gguf_free(r)


# ------------------------------------------------------------------------------------------------ GPU: the loader
@pytest.mark.gpu
def test_device_image_equals_file_bytes(tmp_path, kern):
    """small tensors sharing a stage, tensors larger than one 16 MiB stage, sizes that are not multiples of anything"""
    tensors = gf.sample_tensors(seed=5, big=True)          # 46 MB + 30 MB tensors among the small ones
    p = str(tmp_path / "w.gguf")
    gf.write(p, gf.sample_kvs(), tensors)
    with b200.Gguf(p, kern.lib) as g:
        for rep in range(2):                                # the staging ring is reused by the second load
            buf = g.load(kern)
            kern.torch.cuda.synchronize()
            host = buf.cpu().numpy()
            for info, (_, _, _, raw) in zip(g.tensors(), tensors):
                got = host[info["device_offset"]:info["device_offset"] + info["nbytes"]].tobytes()
                assert got == raw, (info["name"], rep)
        # chosen tensors at chosen addresses
        infos = g.tensors()
        ids = [g.find_tensor("big.q6_K"), g.find_tensor("one"), g.find_tensor("blk.0.attn_q.weight")]
        dsts = [kern.torch.zeros(infos[i]["nbytes"] + 64, dtype=kern.torch.uint8, device=kern.device) for i in ids]
        g.load_tensors(kern, ids, [d.data_ptr() + 16 for d in dsts])
        kern.torch.cuda.synchronize()
        for i, d in zip(ids, dsts):
            h = d.cpu().numpy()
            assert h[16:16 + infos[i]["nbytes"]].tobytes() == tensors[i][3]
            assert not h[:16].any() and not h[16 + infos[i]["nbytes"]:].any()
        # argument errors
        assert kern.lib.b200_gguf_load(g.h, buf.data_ptr(), 10, None) == 4           # B200_ERR_WORKSPACE
        assert kern.lib.b200_gguf_load(g.h, buf.data_ptr() + 8, buf.numel(), None) == 3   # B200_ERR_STRIDE


@pytest.mark.gpu
def test_loaded_weights_feed_mul_mat(tmp_path, kern, oracle):
    """quantized by the oracle, written to GGUF, streamed to the device, used in place by the decode mat-vec"""
    from oracle_lib import random_blocks
    m, k = 768, 4096
    ws = {t: np.ascontiguousarray(random_blocks(t, m, k, seed=31 + i)) for i, t in enumerate(b200.QUANT_TYPES)}
    tensors = [(f"blk.{i}.w_{t}", [k, m], b200.TYPE_ID[t], ws[t].tobytes()) for i, t in enumerate(b200.QUANT_TYPES)]
    p = str(tmp_path / "m.gguf")
    gf.write(p, [("general.architecture", gf.STR, "test")], tensors)
    x = np.random.default_rng(9).uniform(-1, 1, size=(1, k)).astype(np.float32)
    xd = kern.to_device(x)
    with b200.Gguf(p, kern.lib) as g:
        buf = g.load(kern)
        for info, t in zip(g.tensors(), b200.QUANT_TYPES):
            assert info["ne"][:2] == [k, m]
            w = buf[info["device_offset"]:info["device_offset"] + info["nbytes"]]
            got = kern.mul_mat(t, w, m, k, xd).cpu().numpy().reshape(1, m)
            assert oracle.nmse(oracle.mul_mat(t, ws[t], m, k, x, mode=1), got) <= 1e-9, t
