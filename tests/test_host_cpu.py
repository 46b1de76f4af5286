"""CPU-only checks of the product's host side: the view algebra / lowering of CUDATensor against
the oracle's restatement of the reference (no device memory is touched: tensors adopt a fake
pointer), and the C ABI: the library loads, exports every symbol include/waffe2cuda.h declares, and
refuses to compute without a B200 (no CPU fallback)."""
import ctypes
import importlib
import itertools
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import loopnest as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def fake_tensor(wf, shape, order="column"):
    n = int(np.prod(shape))
    return wf.CUDATensor(wf.Storage(n * 4, ptr=0x10000), shape, order=order)


def lowered_offsets(t):
    """Flat storage offsets of every visible element, from the lowered wf2cu_view."""
    lv = t.lower()
    shape = [lv.shape[i] for i in range(lv.rank)]
    off = np.full(shape if shape else (1,), lv.offset, dtype=np.int64)
    for ax in range(lv.rank):
        sh = [1] * lv.rank
        sh[ax] = shape[ax]
        off = off + (np.arange(shape[ax], dtype=np.int64) * lv.stride[ax]).reshape(sh)
    return off


def oracle_offsets(t):
    """Same, from the oracle's mref arithmetic (tensor.lisp:671-693)."""
    n = len(t.vec)
    t.vec[:] = np.arange(n, dtype=np.float32)
    return t.to_numpy().astype(np.int64)


SUBS_2D = [
    (True, True), ([2, 14], [4, 28]), ([0, 20, 2], True), (True, [0, 36, -1]), ([1, 19, -3], [2, 34, 4]),
    (3, [8, 16]), ([5, 6], True), ([0, -1], [1, -2]), (True, [36, 0, 2]), ([19, 1, 3], [0, 36, 5]),
]


@pytest.mark.parametrize("subs", SUBS_2D)
@pytest.mark.parametrize("order", ["column", "row"])
def test_view_lowering_matches_oracle(wf, subs, order):
    g = wf.view(fake_tensor(wf, (20, 36), order), *subs)
    o = L.view(L.make_tensor([20, 36], order=order), *subs)
    assert g.shape == o.shape
    np.testing.assert_array_equal(lowered_offsets(g), oracle_offsets(o))


def test_view_of_view_and_broadcast(wf):
    g0, o0 = fake_tensor(wf, (10, 12)), L.make_tensor([10, 12])
    chain = [([1, 9], [2, 12]), ([1, 5], [0, 6, 2]), (True, [0, 2])]
    g, o = g0, o0
    for s in chain:
        g, o = wf.view(g, *s), L.view(o, *s)
        assert g.shape == o.shape
        np.testing.assert_array_equal(lowered_offsets(g), oracle_offsets(o))
    gb = wf.view(fake_tensor(wf, (7, 1)), True, [":broadcast", 5])
    ob = L.view(L.make_tensor([7, 1]), True, [":broadcast", 5])
    assert gb.shape == ob.shape == [7, 5]
    np.testing.assert_array_equal(lowered_offsets(gb), oracle_offsets(ob))
    lv = gb.lower()
    assert lv.stride[1] == 0 and lv.stride[0] == 1


@pytest.mark.parametrize("perm", list(itertools.permutations([2, 1, 0])))
def test_permute_lowering_matches_oracle(wf, perm):
    g = wf.permute(fake_tensor(wf, (3, 4, 5)), *perm)
    o = L.permute(L.make_tensor([3, 4, 5]), *perm)
    assert g.shape == o.shape
    np.testing.assert_array_equal(lowered_offsets(g), oracle_offsets(o))


def test_transpose_flexible_permute(wf):
    g = wf.transpose(fake_tensor(wf, (2, 3, 4, 5)))
    o = L.transpose(L.make_tensor([2, 3, 4, 5]))
    assert g.shape == o.shape == [2, 3, 5, 4]
    np.testing.assert_array_equal(lowered_offsets(g), oracle_offsets(o))
    with pytest.raises(ValueError):
        wf.permute(fake_tensor(wf, (2, 3)), 0, 0)


def test_range_algebra_matches_oracle(wf):
    for f, t, s in itertools.product(range(0, 7), range(0, 7), (-3, -2, -1, 1, 2, 3)):
        if f == t:
            continue
        r = (f, t, s)
        assert wf.range_size(r) == L.range_size(L.make_range(*r))
        for i in range(3):
            assert wf.range_nth(r, i) == L.range_nth(L.make_range(*r), i)


def test_strides(wf):
    assert wf.calc_strides([2, 3, 4], "column") == [12, 4, 1]
    assert wf.calc_strides([2, 3, 4], "row") == [1, 2, 6]


def test_reshape_requires_contiguous(wf):
    t = fake_tensor(wf, (4, 6))
    assert wf.reshape_contiguous(t, (3, 8)).shape == [3, 8]
    assert wf.reshape_contiguous(t, (2, -1)).shape == [2, 12]
    with pytest.raises(ValueError):
        wf.reshape_contiguous(wf.view(t, [0, 2], [0, 3]), (6,))
    with pytest.raises(ValueError):
        wf.reshape_contiguous(t, (5, 5))


# ------------------------------------------------------------------ the C ABI
def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "waffe2cuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(wf2cu_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol(wf):
    lib = wf._lib.load()
    names = declared_symbols()
    assert len(names) >= 45
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/waffe2cuda.h but not exported"
    # and every prototype the Python binding uses is declared in the header
    assert set(wf._lib.PROTOTYPES) | {"wf2cu_last_error"} <= set(names)


def test_lisp_host_binds_every_declared_symbol():
    """cl-waffe2_b200/lisp/cffi.lisp (the CFFI side a cl-waffe2 maintainer loads) has a `defwf2` / `defcfun` for every
    entry point of the header.  (No Common Lisp in this image: the file is checked by reading it.)"""
    lisp = open(os.path.join(ROOT, "cl-waffe2_b200", "lisp", "cffi.lisp")).read()
    bound = {"wf2cu_" + m.replace("-", "_") for m in re.findall(r"\(defwf2 ([a-z0-9-]+)", lisp)}
    bound |= set(re.findall(r'\(defcfun \("(wf2cu_[a-z0-9_]+)"', lisp))
    missing = [n for n in declared_symbols() if n not in bound]
    assert not missing, missing


def test_view_struct_layout_matches_header(wf):
    V = wf._lib.View
    assert ctypes.sizeof(V) == 8 + 8 + 4 + 4 + 8 * 8 + 8 * 8
    assert V.shape.offset == 24 and V.stride.offset == 24 + 64


def test_no_cpu_fallback_without_device(wf):
    """Without a B200 every compute entry point must fail loudly, never compute on the host."""
    lib = wf._lib.load()
    n = ctypes.c_int(0)
    lib.wf2cu_device_count(ctypes.byref(n))
    if n.value > 0:
        pytest.skip("a CUDA device is visible; this check is for the CPU-only container")
    with pytest.raises(wf.Wf2CudaError):
        wf.init(0)
    buf = (ctypes.c_float * 16)()
    v = wf._lib.make_view(ctypes.addressof(buf), 0, [4, 4], [4, 1])
    assert lib.wf2cu_ew_binary(0, ctypes.byref(v), ctypes.byref(v)) == 4      # WF2CU_ERR_NOT_INITIALISED
    assert b"no CPU fallback" in lib.wf2cu_last_error()
    assert lib.wf2cu_gemm(0, 0, 4, 4, 4, buf, 0, 4, buf, 0, 4, buf, 0, 4, 1, 0, 0, 0) == 4
    assert list(buf) == [0.0] * 16


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under cl-waffe2_b200/ may reference it."""
    pkg = os.path.join(ROOT, "cl-waffe2_b200")
    for dirpath, _, files in os.walk(pkg):
        if "_obj" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".lisp")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "oracle/" not in src and "libwaffe2_port" not in src and "libwaffe2simd_ref" not in src, f


def test_state_dict_key_convention(wf):
    """`{prefix}:{composite}.{nth}.{slot}` (vm/generic-tensor/state-dicts.lisp:71-96,152-235): nth counts
    duplicates from 0; keys parse back with parse-state-dict-key."""
    import importlib
    sd = importlib.import_module("cl-waffe2_b200.state_dict")

    class Opt:
        def state_for(self, idx):
            return [("lr", "single-float", 0.5), ("n", "bit", 0)]
    Opt.__name__ = "Adam"
    named = [("LinearLayer", "weights", "W0"), ("LinearLayer", "bias", "b0"), ("Conv2D", "weight", "C0"),
             ("LinearLayer", "weights", "W1"), ("LinearLayer", "bias", "b1")]
    d = sd.make_state_dict(named, Opt(), optimizers=True)
    assert d["param:linearlayer.0.weights"] == "W0" and d["param:linearlayer.1.weights"] == "W1"
    assert d["param:conv2d.0.weight"] == "C0" and d["param:linearlayer.1.bias"] == "b1"
    assert d["optimizer:linearlayer.0.bias.adam.single-float.lr"] == 0.5
    assert d["optimizer:linearlayer.1.weights.adam.bit.n"] == 0
    assert sd.parse_state_dict_key("param:linearlayer.0.bias") == ("param", ["linearlayer", "0", "bias"])
    assert sd.parse_state_dict_key("optimizer:linearlayer.0.bias.adam.single-float.lr")[0] == "optimizer"
    with pytest.raises(ValueError):
        sd.parse_state_dict_key("weights:linearlayer.0.bias")
    assert len(sd.make_state_dict(named, Opt(), weights=False, optimizers=True)) == 10


# ------------------------------------------------------------------ the Lisp host files (no Common Lisp here: lint by reading)
LISP_DIR = os.path.join(ROOT, "cl-waffe2_b200", "lisp")


def _strip_lisp(src):
    """Source with comments, strings and character literals blanked out."""
    out, i, n = [], 0, len(src)
    while i < n:
        c = src[i]
        if c == ";":
            while i < n and src[i] != "\n":
                i += 1
        elif c == "#" and src[i + 1:i + 2] == "|":
            j = src.index("|#", i)
            out.append("\n" * src[i:j].count("\n"))
            i = j + 2
        elif c == '"':
            i += 1
            while src[i] != '"':
                if src[i] == "\\":
                    i += 1
                if src[i] == "\n":
                    out.append("\n")
                i += 1
            i += 1
            out.append('""')
        elif c == "#" and src[i + 1:i + 2] == "\\":
            i += 3
            out.append("x")
        else:
            out.append(c)
            i += 1
    return "".join(out)


def test_lisp_host_forms_are_balanced():
    files = sorted(f for f in os.listdir(LISP_DIR) if f.endswith(".lisp"))
    assert {"package.lisp", "cffi.lisp", "tensor.lisp", "lowering.lisp", "arithmetic.lisp", "mathematics.lisp", "logical.lisp",
            "matrix-ops.lisp", "fusion.lisp", "im2col.lisp"} <= set(files)
    for f in files:
        depth, line = 0, 1
        for ch in _strip_lisp(open(os.path.join(LISP_DIR, f)).read()):
            if ch == "\n":
                line += 1
            elif ch == "(":
                depth += 1
            elif ch == ")":
                depth -= 1
                assert depth >= 0, f"{f}:{line}: unmatched ')'"
        assert depth == 0, f"{f}: {depth} unclosed '('"


def test_lisp_host_names_exist_in_reference():
    """Every node class the backend implements and every package-qualified symbol it uses is defined by the reference
    (a typo here would only show up when a maintainer loads the files).  Needs /root/reference, which exists in the
    authoring container only."""
    ref = "/root/reference/source"
    if not os.path.isdir(ref):
        pytest.skip("reference sources are not on this machine")
    corpus = []
    for dirpath, _, fs in os.walk("/root/reference"):
        if "/third_party" in dirpath or "/.git" in dirpath:
            continue
        corpus += [open(os.path.join(dirpath, f), errors="replace").read().lower() for f in fs if f.endswith((".lisp", ".asd"))]
    corpus = "\n".join(corpus)
    own = "\n".join(_strip_lisp(open(os.path.join(LISP_DIR, f)).read()) for f in sorted(os.listdir(LISP_DIR)) if f.endswith(".lisp"))
    nodes = set(re.findall(r"\(define-(?:arith-impl|math-impl|rowred)\s+([A-Za-z0-9>\-]+)", own))
    nodes |= set(re.findall(r"\(define-impl(?:-op)?\s+\(([A-Za-z0-9>\-]+)\s+:device", own))
    defined_here = set(re.findall(r"\(defnode\s+\(([A-Za-z0-9>\-]+)", own))
    assert len(nodes) >= 30, sorted(nodes)
    for nd in sorted(nodes - defined_here):
        assert re.search(r"\((?:defnode|define-op)\s+\(" + re.escape(nd.lower()) + r"[\s)]", corpus) or \
            re.search(r"\(define-(?:arith|math|scalar-mat|logical)[a-z-]*\s.*" + re.escape(nd.lower()), corpus) or \
            nd.lower() in corpus, f"node {nd} is not defined by the reference"
    qualified = set(re.findall(r"\b((?:cl-waffe2/[a-z.\-]+|wf/[a-z]+)::?[a-zA-Z0-9*+<>=/!?\-]+)", own))
    assert qualified, "expected package-qualified symbols"
    for q in sorted(qualified):
        pkg, sym = re.split(r"::?", q, maxsplit=1)
        if pkg.lower() == "cl-waffe2/backends.cuda":
            # our own package, named from the example-side file: the symbol must be defined or bound by the backend's files
            structs = re.findall(r"\(defstruct \(?([a-z0-9-]+)", own, re.I)
            accessor = any(sym.lower().startswith(st.lower() + "-") and
                           re.search(r"\(" + re.escape(sym[len(st) + 1:]) + r"\s", own, re.I) for st in structs)
            assert accessor or re.search(r"\((?:defun|defmacro|defwf2|defclass|defvar|defparameter)\s+\(?" + re.escape(sym) + r"[\s)]",
                                         own, re.I), f"{q} is not defined in lisp/"
            continue
        assert pkg.lower() in corpus, f"package {pkg} ({q}) not found in the reference"
        if sym.lower().endswith("-backward"):
            # `define-op` names the backward node (symb name '-backward) in the defining package (vm/nodes/static-node.lisp:236,272)
            stem = sym.lower()[:-len("-backward")]
            assert re.search(r"\(define-op\s+\(" + re.escape(stem) + r"[\s)]", corpus), f"{q}: no (define-op ({stem} ...)) in the reference"
            continue
        assert re.search(r"[\s(:'#]" + re.escape(sym.lower()) + r"[\s)]", corpus), f"{q}: {sym} not found in the reference"


# ------------------------------------------------------------------ property test: random view / permute chains
def _random_subscript(draw, st, size):
    """One subscript the reference's parse-view accepts for an axis of SIZE elements (view.lisp:138-206)."""
    kind = draw(st.sampled_from(["t", "index", "slice", "step", "neg"] if size > 1 else ["t", "index", "bcast"]))
    if kind == "t":
        return True
    if kind == "index":
        return draw(st.integers(0, size - 1))
    if kind == "bcast":
        return [":broadcast", draw(st.integers(1, 4))]
    a = draw(st.integers(0, size - 1))
    b = draw(st.integers(a + 1, size))
    if kind == "slice":
        return [a, b]
    step = draw(st.integers(1, 3))
    return [a, b, step] if kind == "step" else [b, a, -step] if a > 0 or b < size else [a, b, step]


def test_random_view_chains_match_oracle(wf):
    """Shapes of rank 1-4, up to three views stacked on each other (index / slice / stepped / reversed / broadcast),
    with a permute in between: the lowered (offset, shape, stride) of the product addresses exactly the elements the
    oracle's restatement of `mref` + `range-nth` does; where the reference's algebra leaves the storage (an index into a
    broadcast axis; `permute*` re-applying a slice that does not start at 0: detach-and-clone1, tensor.lisp:880-896) the
    product refuses to lower the view instead of handing a kernel an out-of-range descriptor."""
    hyp = pytest.importorskip("hypothesis")
    st = hyp.strategies

    def same(g, o, msg):
        """True when both address the same elements; False when both leave the storage."""
        assert g.shape == o.shape, msg
        try:
            want = oracle_offsets(o)
        except IndexError:
            with pytest.raises(ValueError, match="outside its storage"):
                g.lower()
            return False
        np.testing.assert_array_equal(lowered_offsets(g), want, err_msg=msg)
        return True

    @hyp.settings(max_examples=int(os.environ.get("WF2_HYP_EXAMPLES", "300")), deadline=None, derandomize="WF2_HYP_EXAMPLES" not in os.environ,
                  suppress_health_check=list(hyp.HealthCheck), database=None)
    @hyp.given(st.data())
    def run(data):
        draw = data.draw
        shape = draw(st.lists(st.integers(1, 6), min_size=1, max_size=4))
        order = draw(st.sampled_from(["column", "row"]))
        g, o = fake_tensor(wf, tuple(shape), order), L.make_tensor(list(shape), order=order)
        for depth in range(draw(st.integers(1, 3))):
            subs = [_random_subscript(draw, st, n) for n in g.shape]
            try:
                o2 = L.view(o, *subs)
            except (AssertionError, ValueError, IndexError, ZeroDivisionError):
                # a chain the reference's view algebra rejects must be rejected by the product as well
                with pytest.raises((AssertionError, ValueError, IndexError, ZeroDivisionError)):
                    wf.view(g, *subs)
                return
            g, o = wf.view(g, *subs), o2
            if 0 in o.shape:            # a stepped range can come out empty (range-size floors): nothing left to view
                assert g.shape == o.shape
                return
            if not same(g, o, f"{shape} {order} {subs}"):
                return
            if depth == 0 and len(g.shape) > 1 and draw(st.booleans()):
                perm = draw(st.permutations(list(range(len(g.shape)))))
                g, o = wf.permute(g, *perm), L.permute(o, *perm)
                if not same(g, o, f"{shape} {order} perm {perm}"):
                    return

    run()


def _c_prototypes():
    """name -> [CFFI type keyword per parameter] parsed from include/waffe2cuda.h."""
    hdr = open(os.path.join(ROOT, "include", "waffe2cuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", " ", hdr, flags=re.S)
    protos = {}
    for ret, name, params in re.findall(r"\b(int|void|const char\s*\*)\s+(wf2cu_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", hdr):
        types = []
        for prm in [p.strip() for p in params.split(",") if p.strip() and p.strip() != "void"]:
            if "*" in prm or "[" in prm:
                types.append(":pointer")
                continue
            words = prm.replace("const", " ").split()
            ctype = " ".join(words[:-1]) if len(words) > 1 else words[0]
            types.append({"int": ":int", "int64_t": ":int64", "uint64_t": ":uint64", "float": ":float", "double": ":double",
                          "size_t": ":size", "unsigned": ":uint", "unsigned int": ":uint", "uint32_t": ":uint32"}[ctype])
        protos[name] = types
    return protos


def test_lisp_bindings_match_header_arity_and_types():
    """Every `defwf2` / `defcfun` in lisp/cffi.lisp passes the number of arguments the C prototype takes, of CFFI types
    that agree with the C parameter types position by position (no Common Lisp here to find a mismatch by calling)."""
    protos = _c_prototypes()
    assert len(protos) >= 75
    lisp = _strip_lisp(open(os.path.join(LISP_DIR, "cffi.lisp")).read())
    checked = 0
    for m in re.finditer(r"\(defwf2 ([a-z0-9-]+)\s*\(", lisp):
        name = "wf2cu_" + m.group(1).replace("-", "_")
        # the parameter list: balanced parens starting at the '(' that ends the match
        i = depth = m.end() - 1
        depth = 0
        while True:
            depth += lisp[i] == "("
            depth -= lisp[i] == ")"
            i += 1
            if depth == 0:
                break
        plist = lisp[m.end():i - 1]
        types = re.findall(r"\(\s*[a-z0-9*-]+\s+(:[a-z0-9]+|\(:struct [a-z0-9-]+\))\s*\)", plist)
        n_params = len(re.findall(r"\(\s*[a-z0-9*-]+\s+", plist))
        assert n_params == len(types), f"{name}: could not read every parameter type in {plist!r}"
        assert name in protos, f"{name} is bound in cffi.lisp but not declared in the header"
        assert types == protos[name], f"{name}: cffi.lisp passes {types}, the header declares {protos[name]}"
        checked += 1
    assert checked >= 70, checked
