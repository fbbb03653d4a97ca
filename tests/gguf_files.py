"""TEST HARNESS: a minimal GGUF writer (format: reference include/gguf.h:1-31) for fixtures, valid and deliberately broken.

Pinned in tests/test_gguf.py: files written here are accepted by the reference's own gguf_init_from_file (oracle/_ref) with the
same keys / tensors, and the broken ones are refused by it too.
"""
from __future__ import annotations

import struct

import numpy as np

U8, I8, U16, I16, U32, I32, F32, BOOL, STR, ARR, U64, I64, F64 = range(13)
_FMT = {U8: "<B", I8: "<b", U16: "<H", I16: "<h", U32: "<I", I32: "<i", F32: "<f", BOOL: "<b", U64: "<Q", I64: "<q", F64: "<d"}


def gstr(s) -> bytes:
    b = s if isinstance(s, bytes) else s.encode()
    return struct.pack("<Q", len(b)) + b


def kv_bytes(key, vtype, value, arr_type=None) -> bytes:
    out = gstr(key) + struct.pack("<i", vtype)
    if vtype == ARR:
        out += struct.pack("<i", arr_type) + struct.pack("<Q", len(value))
        for v in value:
            out += gstr(v) if arr_type == STR else struct.pack(_FMT[arr_type], v)
    elif vtype == STR:
        out += gstr(value)
    else:
        out += struct.pack(_FMT[vtype], value)
    return out


def tensor_info_bytes(name, ne, type_id, offset) -> bytes:
    out = gstr(name) + struct.pack("<I", len(ne))
    for d in ne:
        out += struct.pack("<q", d)
    return out + struct.pack("<i", type_id) + struct.pack("<Q", offset)


def pad(n: int, a: int) -> int:
    return (n + a - 1) // a * a


def build(kvs, tensors, alignment=32, version=3, magic=b"GGUF", n_tensors=None, n_kv=None, offsets=None, data=True) -> bytes:
    """kvs: [(key, type, value[, arr_type])]; tensors: [(name, ne, ggml type id, raw bytes)] -> the file's bytes.
    `general.alignment` is added when alignment != 32.  offsets: override the tensor data offsets (broken files)."""
    kvs = list(kvs)
    if alignment != 32 and not any(k[0] == "general.alignment" for k in kvs):
        kvs.append(("general.alignment", U32, alignment))
    head = magic + struct.pack("<I", version) + struct.pack("<q", len(tensors) if n_tensors is None else n_tensors)
    head += struct.pack("<q", len(kvs) if n_kv is None else n_kv)
    for kv in kvs:
        head += kv_bytes(*kv)
    a = alignment if alignment > 0 and alignment & (alignment - 1) == 0 else 32
    off, blob = 0, b""
    for i, (name, ne, type_id, raw) in enumerate(tensors):
        head += tensor_info_bytes(name, ne, type_id, off if offsets is None else offsets[i])
        raw = bytes(raw)
        blob += raw + b"\0" * (pad(len(raw), a) - len(raw))
        off += pad(len(raw), a)
    head += b"\0" * (pad(len(head), a) - len(head))
    return head + (blob if data else b"")


def write(path, *args, **kw) -> bytes:
    b = build(*args, **kw)
    with open(path, "wb") as f:
        f.write(b)
    return b


def sample_kvs():
    return [
        ("general.architecture", STR, "llama"),
        ("general.name", STR, "b200 fixture é"),
        ("llama.block_count", U32, 2),
        ("llama.context_length", U64, 1 << 33),
        ("llama.rope.freq_base", F32, 10000.0),
        ("llama.attention.layer_norm_rms_epsilon", F64, 1e-5),
        ("t.u8", U8, 200), ("t.i8", I8, -100), ("t.u16", U16, 60000), ("t.i16", I16, -30000), ("t.i32", I32, -7),
        ("t.i64", I64, -(1 << 40)), ("t.bool", BOOL, 1),
        ("tokenizer.ggml.tokens", ARR, ["<s>", "</s>", "", "hello world"], STR),
        ("tokenizer.ggml.scores", ARR, [0.0, -1.5, 2.25], F32),
        ("tokenizer.ggml.token_type", ARR, [1, 2, 3, 4, 5], I32),
        ("t.empty", ARR, [], U16),
    ]


def sample_tensors(seed=0, big=False):
    """a few tensors of the hot-path formats with random bytes (block bytes: tests/oracle_lib.random_blocks where values matter)"""
    rng = np.random.default_rng(seed)
    blk = {2: (32, 18), 8: (32, 34), 12: (256, 144), 13: (256, 176), 14: (256, 210), 0: (1, 4), 1: (1, 2), 26: (1, 4)}
    spec = [("token_embd.weight", [256, 40], 12), ("blk.0.attn_q.weight", [512, 64], 2), ("blk.0.attn_norm.weight", [512], 0),
            ("blk.0.ffn_down.weight", [768, 33], 14), ("blk.0.attn_v.weight", [256, 7, 3], 13), ("odd", [3, 5, 2, 2], 1),
            ("blk.1.attn_k.weight", [64, 9], 8), ("one", [1], 26)]
    if big:
        spec += [("big.q4_K", [4096, 20000], 12), ("big.q6_K", [4096, 9000], 14)]
    out = []
    for name, ne, t in spec:
        qk, bs = blk[t]
        n = ne[0] // qk * bs * int(np.prod(ne[1:], dtype=np.int64))
        out.append((name, ne, t, rng.integers(0, 256, size=n, dtype=np.uint8).tobytes()))
    return out
