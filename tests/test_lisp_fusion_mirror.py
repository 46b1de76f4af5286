"""The Lisp host's `on-finalizing-compiling` rewrites (cl-waffe2_b200/lisp/fusion.lisp), exercised through their Python
mirror (tests/lisp_fusion_mirror.py) on instruction lists recorded from the reference's graph builders, plus lints that tie
the mirror to the Lisp source: every matcher exists under the same name in fusion.lisp and is installed by the method."""
import os
import re

import numpy as np
import pytest

from lisp_fusion_mirror import Graph, Tn, on_finalizing_compiling, run

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FUSION = open(os.path.join(ROOT, "cl-waffe2_b200", "lisp", "fusion.lisp")).read()
rng = np.random.default_rng(7)


def leaf(shape, tmp=False, name=None):
    return Tn(shape, tmp=tmp, data=rng.standard_normal(shape), name=name)


def both(build, bw=None):
    """Build the same graph twice (same seed), run one as recorded and one rewritten; return (launches, result) pairs."""
    res = []
    for rewrite in (False, True):
        global rng
        rng = np.random.default_rng(7)
        g, out, extra = build()
        iseq = g.iseq
        if rewrite:
            iseq, _ = on_finalizing_compiling(iseq, None)
        n = run(iseq)
        res.append((n, np.array(out.arr()), extra, iseq))
    return res


def test_zero_fill_and_sum_become_one_reduce():
    for axis in (0, 1):
        def build():
            g = Graph()
            return g, g.sum(leaf((6, 10)), axis), None
        (n0, r0, _, _), (n1, r1, _, iseq) = both(build)
        assert (n0, n1) == (2, 1)                      # zero fill + A+=B  ->  reduce_sum
        np.testing.assert_array_equal(r0, r1)
        assert [i.op[0] for i in iseq if i.op[0] not in ("node",)] == ["identity", "reduce-sum"]


@pytest.mark.parametrize("avoid_overflow", [False, True])
@pytest.mark.parametrize("axis", [0, 1])
def test_softmax_chain_becomes_one_launch(avoid_overflow, axis):
    def build():
        g = Graph()
        return g, g.softmax(leaf((8, 12)), avoid_overflow=avoid_overflow, axis=axis), None
    (n0, r0, _, _), (n1, r1, _, iseq) = both(build)
    assert n0 == (12 if avoid_overflow else 6)         # copies, zero fills, sums, sub, exp, divs as the reference emits them
    assert n1 == 1
    assert [i.op for i in iseq if i.op[0] == "softmax"] == [("softmax", axis, avoid_overflow)]
    np.testing.assert_allclose(r0, r1, rtol=2e-6, atol=1e-7)


def test_softmax_chain_is_left_alone_when_an_intermediate_is_read_elsewhere():
    def build():
        g = Graph()
        x = leaf((4, 6))
        e = g.exp(x)
        sm = g.arith("DivNode", e, g.sum(e, 1))
        other = g.arith("AddNode", e, sm)              # exp(x) is read again: the chain is not closed
        return g, other, None
    (n0, r0, _, _), (n1, r1, _, iseq) = both(build)
    assert not any(i.op[0] == "softmax" for i in iseq)
    np.testing.assert_array_equal(r0, r1)


def test_training_graph_keeps_places_aligned_and_skips_forward_only_rewrites():
    def build():
        g = Graph(train=True)
        x = leaf((5, 7), tmp=True)
        y = g.arith("MulNode", x, leaf((5, 7)))
        return g, g.softmax(y, avoid_overflow=False), g
    (n0, r0, g0, _), (n1, r1, g1, iseq) = both(build)
    np.testing.assert_array_equal(r0, r1)
    assert not any(i.op[0] == "softmax" for i in iseq)  # Exp / Div save operands for backward: chain left as launches
    mul = next(i for i in iseq if i.op == ("binary3", "MulNode"))
    # (p_copy p_y) -> (nil p_copy p_y); y is user-owned and never written: its copy is dropped, the place adopts the storage
    assert mul.sv4bw[0] is None and mul.sv4bw[1] is not None and mul.sv4bw[2] is None
    assert mul.pre and mul.pre[0][0] == "alias"
    np.testing.assert_array_equal(mul.sv4bw[1].root.data, mul.args[1].arr())
    assert mul.pre[0][1].root.data is mul.args[2].root.data
    assert n1 < n0


def test_relu_and_linear_inference():
    def build():
        g = Graph()
        x, w, b = leaf((16, 32), tmp=True), leaf((24, 32)), leaf((1, 24))
        return g, g.relu(g.linear(x, w, b)), None
    (n0, r0, _, _), (n1, r1, _, iseq) = both(build)
    assert (n0, n1) == (5, 1)                          # matmul, bias add, where, copy, mul  ->  one gemm_ex (bias + relu epilogue)
    np.testing.assert_allclose(r0, r1, rtol=1e-6, atol=1e-6)
    op = next(i.op for i in iseq if i.op[0] == "gemm-ex")
    assert op[3] is True and op[4] == 1


def test_relu_alone():
    def build():
        g = Graph()
        return g, g.relu(leaf((9, 9))), None
    (n0, r0, _, _), (n1, r1, _, _) = both(build)
    assert (n0, n1) == (3, 1)
    np.testing.assert_array_equal(r0, r1)


def test_linear_training_forward_saves_nothing_it_can_alias():
    def build():
        g = Graph(train=True)
        x, w, b = leaf((16, 32), tmp=False, name="x"), leaf((24, 32), name="w"), leaf((1, 24), name="b")
        return g, g.linear(x, w, b), g
    (n0, r0, _, _), (n1, r1, _, iseq) = both(build)
    assert (n0, n1) == (4, 1)                          # sv4bw copies of x and w + matmul + bias  ->  one gemm_ex
    np.testing.assert_allclose(r0, r1, rtol=1e-6, atol=1e-6)
    gemm = next(i for i in iseq if i.op[0] == "gemm-ex")
    assert len(gemm.sv4bw) == len(gemm.args) and all(p is None for p in gemm.sv4bw)
    assert [k for _, _, k in gemm.pre] == [0, 1]


@pytest.mark.parametrize("first", [True, False])
def test_gradient_adder_folds_into_the_gemm(first):
    res = []
    for rewrite in (False, True):
        r = np.random.default_rng(3)
        x = Tn((16, 32), data=r.standard_normal((16, 32)), name="x")
        w = Tn((24, 32), tmp=False, data=r.standard_normal((24, 32)), name="w")
        dout = Tn((16, 24), data=r.standard_normal((16, 24)), name="dout")
        gw = Tn((24, 32), tmp=False, data=r.standard_normal((24, 32)), name="w.grad")
        g = Graph()
        g.linear_backward(dout, x, w, gw, gx_needed=True, first=first)
        iseq = g.iseq
        if rewrite:
            _, iseq = on_finalizing_compiling([], iseq)
        res.append((run(iseq), gw.arr().copy(), iseq))
    (n0, g0, _), (n1, g1, iseq) = res
    assert (n0, n1) == (3, 2)                          # dX gemm, dW gemm, adder copy/add  ->  dX gemm, dW gemm into w.grad
    np.testing.assert_allclose(g0, g1, rtol=1e-5, atol=1e-5)
    folded = [i for i in iseq if i.op[0] == "gemm-ex"]
    assert len(folded) == 1 and folded[0].op[5] == (not first) and folded[0].op[7] is True
    assert (folded[0].op[1], folded[0].op[2]) == (True, False)          # dW = dout^T x


# ---------------------------------------------------------------------------------------------------- lints against fusion.lisp
MATCHERS = ["alias-immutable-sv4bw", "fuse-move+arith", "fuse-matmul+bias", "fuse-zero-fill+sum", "fuse-softmax-chain",
            "fuse-relu", "fuse-gemm+relu", "fold-grad-adder", "fuse-unfold-chain"]


def test_every_mirrored_matcher_is_defined_and_installed():
    method = FUSION[FUSION.index("(defmethod on-finalizing-compiling"):]
    method = method[:method.index("\n\n")]
    for name in MATCHERS:
        assert re.search(r"\(defun " + re.escape(name) + r"[\s(]", FUSION), f"fusion.lisp lacks (defun {name} ...)"
        assert re.search(r"\(" + re.escape(name) + r"[\s)]", method), f"{name} is not called by on-finalizing-compiling"
    import lisp_fusion_mirror as m
    for name in MATCHERS[:-1]:
        assert hasattr(m, name.replace("-", "_").replace("+", "_")), name
