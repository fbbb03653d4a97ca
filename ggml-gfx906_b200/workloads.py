"""Named synthetic workloads of the quantized hot path (BASELINE.json `configs`).

A workload is a list of MUL_MAT problems (type, m, k) plus the token count n; nothing here touches a
device.  Algorithmic bytes follow SURVEY.md §8(d): weights once + f32 activations once + f32 output
once (the q8 scratch is an implementation artefact and is not counted).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List

BLOCK_ELEMS = {"q4_0": 32, "q8_0": 32, "q4_K": 256, "q5_K": 256, "q6_K": 256}
BLOCK_BYTES = {"q4_0": 18, "q8_0": 34, "q4_K": 144, "q5_K": 176, "q6_K": 210}


@dataclass(frozen=True)
class Mat:
    name: str
    type: str
    m: int      # weight rows = output features
    k: int      # reduction length

    @property
    def weight_bytes(self) -> int:
        return self.m * (self.k // BLOCK_ELEMS[self.type]) * BLOCK_BYTES[self.type]

    def algorithmic_bytes(self, n: int) -> int:
        return self.weight_bytes + n * self.k * 4 + n * self.m * 4

    def flops(self, n: int) -> int:
        return 2 * self.m * self.k * n


def llama2_7b_stack(wtype: str = "q4_K", n_layers: int = 32, row_mult: int = 1) -> List[Mat]:
    """BASELINE configs[2]: every projection of Llama-2-7B (32 x {wq,wk,wv,wo 4096x4096, gate/up
    11008x4096, down 4096x11008} + output 32000x4096), all in one format.  3.716 GB as Q4_K.
    row_mult scales m (the row-split dimension) for the weak-scaling multi-GPU runs."""
    mats = []
    for l in range(n_layers):
        for nm in ("wq", "wk", "wv", "wo"):
            mats.append(Mat(f"l{l}.{nm}", wtype, 4096 * row_mult, 4096))
        mats.append(Mat(f"l{l}.gate", wtype, 11008 * row_mult, 4096))
        mats.append(Mat(f"l{l}.up", wtype, 11008 * row_mult, 4096))
        mats.append(Mat(f"l{l}.down", wtype, 4096 * row_mult, 11008))
    mats.append(Mat("output", wtype, 32000 * row_mult, 4096))
    return mats


def llama3_8b_q4_k_m_stack(n_layers: int = 32) -> List[Mat]:
    """BASELINE configs[3]: Llama-3-8B with a Q4_K_M-like mix: wv and down in
    Q6_K, the rest Q4_K, output Q6_K."""
    mats = []
    for l in range(n_layers):
        mats += [Mat(f"l{l}.wq", "q4_K", 4096, 4096), Mat(f"l{l}.wk", "q4_K", 1024, 4096),
                 Mat(f"l{l}.wv", "q6_K", 1024, 4096), Mat(f"l{l}.wo", "q4_K", 4096, 4096),
                 Mat(f"l{l}.gate", "q4_K", 14336, 4096), Mat(f"l{l}.up", "q4_K", 14336, 4096),
                 Mat(f"l{l}.down", "q6_K", 4096, 14336)]
    mats.append(Mat("output", "q6_K", 128256, 4096))
    return mats


def prefill_ffn_shapes() -> List[Mat]:
    """BASELINE configs[1]: Q8_0 and Q4_0, n=512, Llama-2-7B FFN shapes."""
    return [Mat(f"{t}.{m}x{k}", t, m, k) for t in ("q8_0", "q4_0") for (m, k) in ((11008, 4096), (4096, 11008))]


def llama3_70b_ffn(n_layers: int = 1) -> List[Mat]:
    """BASELINE configs[4]: Llama-3-70B FFN in Q4_K, the row-split shapes."""
    mats = []
    for l in range(n_layers):
        mats += [Mat(f"l{l}.gate", "q4_K", 28672, 8192), Mat(f"l{l}.up", "q4_K", 28672, 8192),
                 Mat(f"l{l}.down", "q4_K", 8192, 28672)]
    return mats


def split_rows(m: int, world: int, rank: int, rounding: int = 128):
    """Row range of `rank` for an even tensor_split: boundaries rounded down to `rounding` rows (the
    reference rounds to mmq_y = 128, src/ggml-cuda/ggml-cuda.cu:765-791), the last rank takes the rest."""
    def low(r):
        if r >= world:
            return m
        v = (m * r) // world
        return v - v % rounding
    return low(rank), low(rank + 1)
