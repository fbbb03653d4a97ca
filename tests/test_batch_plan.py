"""CPU tests of the list entry point's host logic (no device): how b200_mul_mat_batch cuts a list into launches and
which node each node waits for inside a persistent launch (b200_mul_mat_batch_plan = the same code in a dry run).
Pointers are fake: the plan only compares address ranges."""
import pytest

import b200

X0, Y0, W0 = 0x1000_0000, 0x2000_0000, 0x4000_0000


@pytest.fixture(scope="module")
def lib():
    return b200.load_library()


def plan(lib, nodes):
    arr = (b200.MulMatNode * len(nodes))(*nodes)
    return b200.batch_plan(lib, arr)


class Arena:
    """named f32 vectors at distinct fake addresses"""
    def __init__(self, base):
        self.next, self.at = base, {}

    def __call__(self, name, n=1 << 16):
        if name not in self.at:
            self.at[name] = self.next
            self.next += 4 * n
        return self.at[name]


def decoder(types, layers=2, hidden=4096, ffn=11008, vocab=32000):
    tq, tk, tv, to, tg, tu, td, tout = types
    buf = Arena(Y0)
    nodes, names = [], []
    cur = "h0"
    for l in range(layers):
        for nm, t in (("q", tq), ("k", tk), ("v", tv)):
            nodes.append((t, hidden, hidden, cur, f"{nm}{l}"))
        nodes += [(to, hidden, hidden, f"q{l}", f"o{l}"), (tg, ffn, hidden, f"o{l}", f"g{l}"),
                  (tu, ffn, hidden, f"o{l}", f"u{l}"), (td, hidden, ffn, f"u{l}", f"d{l}")]
        cur = f"d{l}"
    nodes.append((tout, vocab, hidden, cur, "logits"))
    out = [b200.mul_mat_node(t, W0 + (i << 28), m, k, buf(src), 1, buf(dst), const_weights=True) for i, (t, m, k, src, dst) in enumerate(nodes)]
    return out, [n[4] for n in nodes]


def test_decoder_pass_is_one_launch_with_the_real_dependencies(lib):
    nodes, names = decoder(("q4_K",) * 8)
    p = plan(lib, nodes)
    assert {g for g, _, _ in p} == {0} and {r for _, _, r in p} == {1}
    ix = {nm: i for i, nm in enumerate(names)}
    dep = {nm: (names[d] if d >= 0 else None) for nm, (_, d, _) in zip(names, p)}
    assert dep["q0"] is None and dep["k0"] is None and dep["v0"] is None
    assert dep["o0"] == "q0"
    assert dep["g0"] == "o0" and dep["u0"] is None        # u0 reads o0 too, but g0 already waited for it (waits are in order)
    assert dep["d0"] == "u0"
    assert dep["q1"] == "d0" and dep["k1"] is None and dep["v1"] is None
    assert dep["logits"] == "d1"
    assert all(d < i for i, (_, d, _) in enumerate(p))
    assert ix["o1"] > ix["q1"]


def test_chaining_switched_off_cuts_at_every_link(lib):
    nodes, names = decoder(("q4_K",) * 8, layers=2)
    prev = lib.b200_set_option(b200.OPT["no_chain"], 1)
    try:
        p = plan(lib, nodes)
    finally:
        lib.b200_set_option(b200.OPT["no_chain"], prev)
    groups = [g for g, _, _ in p]
    assert all(d == -1 for _, d, _ in p)
    # q k v | o | g u | d  per layer, then the output matrix
    assert groups == [0, 0, 0, 1, 2, 2, 3, 4, 4, 4, 5, 6, 6, 7, 8]


def test_mixed_format_families_chain_inside_a_family_and_cut_between(lib):
    # Q4_K_M-like: wv and down (and the output matrix) in Q6_K = another kernel instantiation = another launch
    nodes, names = decoder(("q4_K", "q4_K", "q6_K", "q4_K", "q4_K", "q4_K", "q6_K", "q6_K"), layers=2)
    p = dict(zip(names, plan(lib, nodes)))
    assert p["q0"][0] == p["k0"][0] == p["v0"][0] == p["o0"][0] == p["g0"][0] == p["u0"][0]   # v0 (Q6_K) has no hazard: same flush
    assert p["o0"][1] == names.index("q0") and p["g0"][1] == names.index("o0")
    assert p["d0"][0] == p["u0"][0] + 1 and p["d0"][1] == -1          # reads a Q4_K node's output from the Q6_K launch: cut
    assert p["q1"][0] == p["d0"][0] + 1                              # and back
    assert p["logits"][0] == p["d1"][0] and p["logits"][1] == names.index("d1")   # Q6_K reads Q6_K: chained


def test_write_after_read_and_write_after_write_become_dependencies(lib):
    buf = Arena(Y0)
    k = 1024
    mk = lambda i, src, dst: b200.mul_mat_node("q4_K", W0 + (i << 24), k, k, buf(src), 1, buf(dst))   # noqa: E731
    nodes = [mk(0, "x", "A"), mk(1, "A", "B"), mk(2, "B", "A"), mk(3, "A", "C"), mk(4, "x", "C"), mk(5, "C", "B")]
    p = plan(lib, nodes)
    assert [g for g, _, _ in p] == [0] * 6
    assert [d for _, d, _ in p] == [-1, 0, 1, 2, 3, 4]
    # the third node overwrites what the first wrote and the second read; the fifth overwrites C (written by the fourth)


def test_independent_nodes_have_no_dependencies_and_share_the_launch(lib):
    xs, ys = Arena(X0), Arena(Y0)
    nodes = [b200.mul_mat_node("q4_K", W0 + (i << 26), 4096, 4096, xs(f"x{i}"), 1, ys(f"y{i}")) for i in range(40)]
    assert plan(lib, nodes) == [(0, -1, 1)] * 40
    # more nodes than one launch holds (448): the list is cut, nothing else changes
    nodes = [b200.mul_mat_node("q4_K", W0 + (i << 20), 64, 1024, xs(f"x{i}", 1024), 1, ys(f"y{i}", 64)) for i in range(460)]
    p = plan(lib, nodes)
    assert [g for g, _, _ in p] == [0] * 448 + [1] * 12


def test_few_column_nodes_take_the_mma_list_and_hazards_cut_it(lib):
    xs, ys = Arena(X0), Arena(Y0)
    a = b200.mul_mat_node("q4_K", W0, 4096, 4096, xs("x"), 4, ys("a"))
    b = b200.mul_mat_node("q4_K", W0 + (1 << 26), 4096, 4096, ys("a"), 4, ys("b"))     # reads a: another launch
    c = b200.mul_mat_node("q8_0", W0 + (2 << 26), 4096, 4096, xs("x"), 4, ys("c"))     # Q8_0: the MMA list too (its own ring family)
    d = b200.mul_mat_node("q6_K", W0 + (3 << 26), 64, 2304, xs("x2"), 4, ys("d"))      # Q6_K, 9 super-blocks (rows at 2-byte alignment): the MMA list
    e = b200.mul_mat_node("q8_0", W0 + (4 << 26) + 8, 64, 4096, xs("x3"), 4, ys("e"))   # weights not 16-byte aligned: one persistent entry per column
    p = plan(lib, [a, b, c, d, e])
    assert p[0] == (0, -1, 2) and p[1] == (1, -1, 2)
    assert p[2][2] == 2 and p[2][1] == -1
    assert p[3][2] == 2 and p[3][1] == -1
    assert p[4][2] == 1 and p[4][1] == -1


def test_shapes_outside_the_list_kernels_go_node_by_node(lib):
    xs, ys = Arena(X0), Arena(Y0)
    small_k = b200.mul_mat_node("q4_K", W0, 64, 512, xs("x"), 1, ys("y"))              # k < 1024
    prefill = b200.mul_mat_node("q4_K", W0 + (1 << 26), 4096, 4096, xs("p"), 512, ys("z", 4096 * 512))
    ok = b200.mul_mat_node("q4_K", W0 + (2 << 26), 4096, 4096, xs("q"), 1, ys("w"))
    p = plan(lib, [ok, small_k, prefill, ok])
    assert [r for _, _, r in p] == [1, 0, 0, 1]
    assert len({g for g, _, _ in p}) == 4
