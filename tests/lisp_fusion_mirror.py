"""Python mirror of cl-waffe2_b200/lisp/fusion.lisp's instruction-sequence rewrites (test infrastructure).

No Common Lisp exists in the authoring image, so the matchers of `on-finalizing-compiling` are restated here function by
function (same names, same conditions, same order) over a small stand-in for the reference's `WfInstruction`
(`source/vm/ir.lisp:13-65`: op, node, args, out-to, sv4bw, grad-adder-p) and run on instruction lists recorded by hand from
the reference's graph-building functions (`!sum` reduction.lisp:60-79, `!softmax` nn/activation.lisp:153-159, `!relu`
activation.lisp:26, LinearLayer nn/regression.lisp:14-18, `expand-gradient-adder` vm/utils.lisp:109-142).  A numpy
interpreter executes an instruction list before and after rewriting: the rewrites must not change any result, and the launch
counts are what INTEGRATION.md quotes.
"""
import numpy as np


class Tn:
    """AbstractTensor stand-in: a view (per axis True or ('broadcast', n)) of a storage root."""
    _count = 0

    def __init__(self, shape, root=None, view=None, tmp=True, data=None, name=None):
        Tn._count += 1
        self.id = Tn._count
        self.shape = tuple(shape)
        self.root = root.root if root is not None else self
        self.view = tuple(view) if view is not None else (True,) * len(self.shape)
        self.tmp = tmp                         # tensor-tmp-p: storage owned by the VM's memory pool
        self.name = name or f"t{self.id}"
        if root is None:
            actual = tuple(1 if v is not True else s for s, v in zip(self.shape, self.view))
            self.data = np.zeros(actual, np.float32) if data is None else np.asarray(data, np.float32).reshape(actual)

    def alias(self):
        """The tensor an in-place node returns: a new object over the same storage and view (`out-to`)."""
        return Tn(self.shape, root=self, view=self.view, tmp=self.tmp, name=self.name + "'")

    def arr(self):
        return np.broadcast_to(self.root.data, self.shape)

    def bcast_axes(self):
        return tuple(k for k, v in enumerate(self.view) if v is not True)

    def __repr__(self):
        return f"<{self.name}{list(self.shape)}>"


class Inst:
    def __init__(self, node, args, out_to, sv4bw=None, attrs=None, grad_adder=False):
        self.node, self.args, self.out_to = node, list(args), list(out_to)
        self.sv4bw = list(sv4bw) if sv4bw else []
        self.attrs = dict(attrs or {})
        self.grad_adder = grad_adder
        self.op = ("node",)                     # what the closure in `wfop-op` does; rewrites replace it
        self.pre = []                           # host-side work a rewrite puts in front of the op (no launch)

    def __repr__(self):
        return f"{self.node}{self.op if self.op != ('node',) else ''}({', '.join(map(repr, self.args))}) -> {self.out_to}"


ARITH = {"AddNode": np.add, "SubNode": np.subtract, "MulNode": np.multiply, "DivNode": np.divide}
NO_LAUNCH = {"ViewTensorNode", "LazyTransposeNode", "Permute-Node"}


# ----------------------------------------------------------------------------------------------------------- interpreter
def _store(dst, value):
    """dst <- value through dst's view; a destination with broadcast axes receives the SUM over them (the reference's
    `A+=B out* x` reduction idiom is handled by the caller, plain stores need a dense destination)."""
    assert not dst.bcast_axes(), "store into a broadcast view"
    dst.root.data[...] = np.asarray(value, np.float32).reshape(dst.root.data.shape)


def untransposed(t):
    """`read-untransposed` (matrix-ops.lisp:52-56): the tensor under a `!t`."""
    return getattr(t, "src", None) or t


def _gemm(a, b, ta, tb):
    """TA / TB as MatMulNode's `trans-a?` / `trans-b?`: the operand arrived through `!t` and is read untransposed."""
    A = untransposed(a).arr().T if ta else a.arr()
    B = untransposed(b).arr().T if tb else b.arr()
    return (A.astype(np.float64) @ B.astype(np.float64)).astype(np.float32)


def run(iseq, no_grad=False):
    """Execute; returns the number of kernel launches (sv4bw copies included, `apply-inst-sv4bw` vm.lisp:40-65)."""
    launches = 0
    for ins in iseq:
        if not no_grad:
            for var, place in zip(ins.args, ins.sv4bw):
                if place is not None and var is not None:
                    place.root.data = var.root.data.copy()
                    launches += 1
        for kind, place, k in ins.pre:
            assert kind == "alias"
            place.root.data = ins.args[k].root.data      # (setf (tensor-vec place) (tensor-vec var)): no copy
        op, a = ins.op, ins.args
        if op[0] == "identity":
            continue
        if op[0] == "node":
            if ins.node in NO_LAUNCH:                   # views / lazy transposes: host pointer work only
                continue
            launches += 1
            if ins.node == "MoveTensorNode":
                _store(a[0], _read(a[1]))
            elif ins.node in ARITH:
                f = ARITH[ins.node]
                if a[0].bcast_axes():                   # reduction idiom: only AddNode reaches here
                    assert ins.node == "AddNode"
                    ax = a[0].bcast_axes()
                    a[0].root.data[...] = a[0].root.data + _read(a[1]).sum(axis=ax, keepdims=True, dtype=np.float32)
                else:
                    _store(a[0], f(a[0].arr(), _read(a[1])))
            elif ins.node == "ExpNode":
                _store(a[1], np.exp(_read(a[0])))
            elif ins.node == "Where-Operation-Node":
                _store(a[1], np.where(_read(a[0]) > ins.attrs["scal"], np.float32(1), np.float32(0)))
            elif ins.node == "MatMulNode":
                _store(a[2], _gemm(a[0], a[1], ins.attrs.get("ta", False), ins.attrs.get("tb", False)))
            else:
                raise NotImplementedError(ins.node)
            continue
        launches += 1
        if op[0] == "binary3":                          # dst <- x op y
            _store(a[0], ARITH[op[1]](_read(a[1]), _read(a[2])))
        elif op[0] == "reduce-sum":                     # out* <- sum over its broadcast axes of x
            ax = a[0].bcast_axes()
            a[0].root.data[...] = _read(a[1]).sum(axis=ax, keepdims=True, dtype=np.float32)
        elif op[0] == "softmax":                        # args (dst, x)
            x = _read(a[1]).astype(np.float32)
            if op[2]:
                x = x - x.mean(axis=op[1], keepdims=True, dtype=np.float32)
            e = np.exp(x)
            _store(a[0], e / e.sum(axis=op[1], keepdims=True, dtype=np.float32))
        elif op[0] == "relu":                           # args (dst, x)
            x = _read(a[1])
            _store(a[0], x * np.where(x > 0, np.float32(1), np.float32(0)))
        elif op[0] == "gemm-ex":                        # args (a, b, out[, bias]); op = (_, ta, tb, bias?, act, accumulate, dst)
            _, ta, tb, has_bias, act, acc, dst, transposed_dst = op
            c = _gemm(a[0], a[1], ta, tb)
            if has_bias:
                c = c + a[3].arr()
            if act == 1:
                c = c * (c > 0)
            if acc:
                c = dst.arr() + c
            _store(dst, c)
        else:
            raise NotImplementedError(op)
    return launches


def _read(t):
    src = getattr(t, "src", None)
    return src.arr().T if src is not None else t.arr()


# ------------------------------------------------------------------------------- recorded instruction lists (builders)
class Graph:
    """Appends instructions in the order `node-compile-into-vm` emits them (topological, producers first:
    `vm/compile.lisp:72-104`).  `train=True` gives every node the save-for-backward places its `defnode` declares."""

    def __init__(self, train=False):
        self.iseq, self.train = [], train

    def emit(self, node, args, out, sv=None, **attrs):
        sv4bw = None
        if self.train and sv:
            sv4bw = [Tn(a.root.data.shape, name="sv." + a.name) if keep else None for a, keep in zip(args, sv)]
        ins = Inst(node, args, [out], sv4bw=sv4bw, attrs=attrs)
        self.iseq.append(ins)
        return out

    def view(self, base, view, shape):                   # `!view` -> ViewTensorNode(result, before), fundamental.lisp:173-184
        v = Tn(shape, root=base, view=view, tmp=base.tmp, name=base.name + "*")
        return self.emit("ViewTensorNode", [v, base], v.alias())

    def copy(self, x, ignore_me=False):                  # `!copy` = Move(make-input, x), fundamental.lisp:9-43
        c = Tn(x.shape, name="copy")
        return self.emit("MoveTensorNode", [c, x], c.alias(), ignore_me=ignore_me)

    def const(self, value, like):                        # `infer-scalar-tensor` + `!view ... (broadcast-to x)`, arithmetic.lisp:31-51,244-250
        c = Tn((1,) * len(like.shape), tmp=False, data=[value], name=f"const{value:g}")
        return self.view(c, tuple(("broadcast", s) for s in like.shape), like.shape)

    def arith(self, node, x, y, in_place=False):         # `!add` & co: (forward (XNode) (!copy x) y), arithmetic.lisp:150-152
        d = x if in_place else self.copy(x)
        if isinstance(y, float):                         # a number: its one-element tensor and window come after the copy
            y = self.const(y, d)
        sv = {"AddNode": None, "SubNode": None, "MulNode": (True, True), "DivNode": (True, True)}[node]
        return self.emit(node, [d, y], d.alias(), sv=sv)

    def sum(self, x, axis, div_by=None):                 # `!sum ... :keepdims t`, reduction.lisp:60-79
        shape = tuple(1 if k == axis else s for k, s in enumerate(x.shape))
        out = Tn(shape, name="sum")
        out = self.arith("MulNode", out, 0.0, in_place=True)                       # A*=scal out 0
        view = tuple(("broadcast", s) if k == axis else True for k, s in enumerate(x.shape))
        star = self.view(out, view, x.shape)
        res = self.arith("AddNode", star, x, in_place=True)                        # A+=B out* x
        if div_by is None:
            return res
        back = self.view(res, (True,) * len(shape), shape)                         # reverser
        q = self.arith("DivNode", back, float(div_by))                             # !div out div-by
        return self.view(q, view, x.shape)

    def mean(self, x, axis):                             # `!mean`, reduction.lisp:81-136
        return self.sum(x, axis, div_by=float(x.shape[axis]))

    def exp(self, x):                                    # (forward (ExpNode) x (!copy x :maybe-in-place t)), mathematics.lisp:117-120
        c = self.copy(x)
        return self.emit("ExpNode", [x, c], c.alias(), sv=(True, False))

    def softmax(self, x, avoid_overflow=True, axis=1):   # nn/activation.lisp:153-159
        if avoid_overflow:
            x1 = self.arith("SubNode", x, self.mean(x, axis))
            e = self.exp(x1)
            return self.arith("DivNode", e, self.sum(e, axis))
        e = self.exp(x)
        return self.arith("DivNode", e, self.sum(e, axis))

    def relu(self, x):                                   # (!mul x (A>scal x 0.0)), activation.lisp:26
        m = Tn(x.shape, name="mask")
        m = self.emit("Where-Operation-Node", [x, m], m.alias(), scal=0.0)
        return self.arith("MulNode", x, m)

    def t(self, x):                                      # `!t` = !permute + LazyTransposeNode (matrix-ops.lisp:64-78)
        xt = Tn(x.shape[::-1], root=x, tmp=x.tmp, name=x.name + ".T")   # same storage, read through `.src`
        xt.src = x
        self.emit("Permute-Node", [x], xt)
        return self.emit("LazyTransposeNode", [xt], xt)

    def linear(self, x, w, b):                           # (!add (!matmul x (!t weight)) (%transform bias[i] -> [~ i])), regression.lisp:14-18
        wt = self.t(w)
        out = Tn((x.shape[0], w.shape[0]), name="mm")
        mm = self.emit("MatMulNode", [x, wt, out], out.alias(), sv=(True, True, False), ta=False, tb=True)
        bias = self.view(b, (("broadcast", x.shape[0]), True), (x.shape[0], b.shape[-1]))
        return self.arith("AddNode", mm, bias, in_place=True)

    def linear_backward(self, dout, x, w, gw, gx_needed, first=True):
        """Backward of MatMulNode(x, (!t w)) (matrix-ops.lisp:10-15) + the gradient adder for w
        (`expand-gradient-adder`, vm/utils.lisp:109-142): dW' = (!matmul (!t x) dout) is (in out); `!t`'s Permute-Node
        backward hands the adder its transposed window."""
        out = []
        if gx_needed:
            gx = Tn(x.shape, name="gx")
            out.append(self.emit("MatMulNode", [dout, w, gx], gx.alias(), ta=False, tb=False))   # dout (!t (!t w)) = dout w
        g = Tn((x.shape[1], dout.shape[1]), name="gwT")
        g1 = self.emit("MatMulNode", [self.t(x), dout, g], g.alias(), ta=True, tb=False)
        gp = Tn(w.shape, name="gw.window")
        gp.src = g1
        self.emit("Permute-Node", [g1], gp)
        adder = Inst("MoveTensorNode" if first else "AddNode", [gw, gp], [gw.alias()], grad_adder=True)
        self.iseq.append(adder)
        return out


# ---------------------------------------------------------------------------------------- the rewrites (fusion.lisp)
LAUNCHING_ARITH = set(ARITH)


def producers(iseq):
    """out-to tensor -> instruction (hash table :test #'eq in fusion.lisp)."""
    return {id(o): ins for ins in iseq for o in ins.out_to}


def use_counts(iseq):
    uses = {}
    for ins in iseq:
        if ins.op == ("identity",):                    # a suppressed instruction reads nothing
            continue
        for a in ins.args:
            uses[id(a)] = uses.get(id(a), 0) + 1
    return uses


def no_sv4bw(ins):
    return all(p is None for p in ins.sv4bw)


def suppress(ins):
    """The instruction keeps its place and bookkeeping but launches nothing: its closure returns the argument that
    `out-to` aliases (first for arithmetic / Move, the OUT operand for `(x out)` nodes)."""
    ins.op = ("identity",)


def zero_constant_p(t, prod):
    """T is (a broadcast view of) a one-element constant 0 made at build time."""
    p = prod.get(id(t))
    base = p.args[1] if p is not None and p.node == "ViewTensorNode" else t
    return (not base.tmp) and base.root.data.size == 1 and float(base.root.data.ravel()[0]) == 0.0


def alias_immutable_sv4bw(iseq_fw, iseq_bw):
    """A save-for-backward copy whose source is user-owned storage (not pool memory) that no instruction of either
    sequence writes is dropped: the place adopts the source's storage when the instruction runs."""
    written = {id(o.root) for ins in iseq_fw + (iseq_bw or []) if ins.node not in NO_LAUNCH for o in ins.out_to}
    for ins in iseq_fw:
        for k, place in enumerate(ins.sv4bw):
            var = ins.args[k]
            if place is not None and not var.tmp and id(var.root) not in written:
                ins.sv4bw[k] = None
                ins.pre.append(("alias", place, k))
    return iseq_fw


def next_launching(iseq, k):
    """Index of the first instruction after K that is neither suppressed nor a window (ViewTensorNode / Permute-Node / LazyTransposeNode:
    the topological order puts the window of a second operand between a producer and its consumer), or None."""
    j = k + 1
    while j < len(iseq) and (iseq[j].node in NO_LAUNCH or iseq[j].op == ("identity",)):
        j += 1
    return j if j < len(iseq) else None


def fuse_move_arith(iseq):
    """copy <- Move(copy, x) ; copy <- Arith(copy, y)  ==>  copy <- binary3(op, copy, x, y); the Arith's places are
    re-aligned to the new argument list: (p_copy p_y) -> (nil p_copy p_y), because `copy` held x when it was saved."""
    out, k = [], 0
    while k < len(iseq):
        i0 = iseq[k]
        j = next_launching(iseq, k)
        i1 = iseq[j] if j is not None else None
        if (i1 is not None and i0.node == "MoveTensorNode" and i0.op == ("node",) and i1.node in ARITH and i1.op == ("node",)
                and not i0.grad_adder and no_sv4bw(i0) and i0.out_to[0] is i1.args[0] and not i1.args[0].bcast_axes()):
            i1.op = ("binary3", i1.node)
            i1.args = [i0.args[0], i0.args[1], i1.args[1]]
            if i1.sv4bw:
                i1.sv4bw = [None, i1.sv4bw[0], i1.sv4bw[1]]
            out += iseq[k + 1:j] + [i1]
            k = j + 1
        else:
            out.append(i0)
            k += 1
    return out


def bias_operand_p(t, n):
    return len(t.shape) == 2 and t.view[0] is not True and t.view[1] is True and t.shape[1] == n


def fuse_matmul_bias(iseq):
    """out <- MatMul(a, b, out) ; out <- Add(out, bias*)  ==>  out <- gemm_ex(a, b, out, bias)"""
    out, k = [], 0
    while k < len(iseq):
        i0 = iseq[k]
        j = next_launching(iseq, k)
        i1 = iseq[j] if j is not None else None
        if (i1 is not None and i0.node == "MatMulNode" and i0.op == ("node",) and i1.node == "AddNode" and i1.op == ("node",)
                and no_sv4bw(i1) and len(i0.args[2].shape) == 2 and i0.out_to[0] is i1.args[0]
                and bias_operand_p(i1.args[1], i0.args[2].shape[1])):
            i1.op = ("gemm-ex", i0.attrs["ta"], i0.attrs["tb"], True, 0, False, i0.args[2], False)
            i1.args = i0.args + [i1.args[1]]
            i1.sv4bw = (i0.sv4bw + [None]) if i0.sv4bw else []
            i1.attrs = dict(i0.attrs)
            i1.node = "MatMulNode"
            out += iseq[k + 1:j] + [i1]
            k = j + 1
        else:
            out.append(i0)
            k += 1
    return out


def fuse_zero_fill_sum(iseq):
    """out <- Mul(out, 0*) ; out* <- View(out*, out) ; out* <- Add(out*, x)  ==>  out* <- reduce_sum(x) (no zero fill).
    Returns (iseq, sums): sums maps the Add's out-to to (x, instruction) for the matchers that follow."""
    prod, uses, sums = producers(iseq), use_counts(iseq), {}
    for ia in iseq:
        if ia.node != "AddNode" or ia.op != ("node",) or not ia.args[0].bcast_axes() or not no_sv4bw(ia):
            continue
        pv = prod.get(id(ia.args[0]))
        if pv is None or pv.node != "ViewTensorNode" or uses.get(id(ia.args[0])) != 1:
            continue
        pz = prod.get(id(pv.args[1]))
        if (pz is None or pz.node != "MulNode" or pz.op != ("node",) or not no_sv4bw(pz) or uses.get(id(pv.args[1])) != 1
                or not zero_constant_p(pz.args[1], prod)):
            continue
        suppress(pz)
        ia.op = ("reduce-sum",)
        sums[id(ia.out_to[0])] = (ia.args[1], ia, pz)
    return iseq, sums


def _through_views(t, prod):
    """Follow ViewTensorNodes back to the tensor they window."""
    while True:
        p = prod.get(id(t))
        if p is None or p.node != "ViewTensorNode":
            return t
        t = p.args[1]


def chain_closed_p(iseq, members, result):
    """MEMBERS (instructions) occupy consecutive positions once view instructions are ignored, none saves anything for
    backward, and nothing they produce is read outside the chain except RESULT."""
    pos = sorted(iseq.index(m) for m in members)
    inside = {id(m) for m in members}
    for p in range(pos[0], pos[-1] + 1):
        if id(iseq[p]) not in inside and iseq[p].node not in NO_LAUNCH:
            return False
    made = {id(o) for m in members for o in m.out_to if o is not result}
    for ins in iseq:                                   # windows of what the chain made belong to the chain
        if ins.node in NO_LAUNCH and any(id(a) in made for a in ins.args):
            made |= {id(o) for o in ins.out_to}
    for ins in iseq:
        if id(ins) not in inside and ins.node not in NO_LAUNCH and any(id(a) in made for a in ins.args):
            return False
    return all(no_sv4bw(m) for m in members)


def fuse_softmax_chain(iseq, sums):
    """The `!softmax` graph (nn/activation.lisp:153-159), after fuse-move+arith and fuse-zero-fill+sum:
         e <- Exp(x1, copy) ; s <- reduce_sum(e) ; out <- binary3(Div, copy2, e, s)
       with x1 = x (avoid-overflow nil) or x1 <- binary3(Sub, copy0, x, mean*) and mean* a window of
       binary3(Div, copy1, reduce_sum(x), n*) over the same axis (avoid-overflow t)
       ==> one softmax launch at the position of the final Div, everything before it suppressed."""
    prod = producers(iseq)
    for idiv in iseq:
        if idiv.op != ("binary3", "DivNode"):
            continue
        dst, e, s = idiv.args
        if id(s) not in sums:
            continue
        sx, isum, izero = sums[id(s)]
        iexp = prod.get(id(e))
        if sx is not e or iexp is None or iexp.node != "ExpNode" or iexp.op != ("node",) or len(s.bcast_axes()) != 1:
            continue
        axis = s.bcast_axes()[0]
        members = [iexp, isum, izero, idiv]
        icopy = prod.get(id(iexp.args[1]))             # the (!copy x) ExpNode writes into
        if icopy is not None and icopy.node == "MoveTensorNode":
            members.append(icopy)
        x1, avoid = iexp.args[0], False
        isub = prod.get(id(x1))
        if isub is not None and isub.op == ("binary3", "SubNode"):
            _, x0, mean_view = isub.args
            imean = prod.get(id(_through_views(mean_view, prod)))
            if imean is not None and imean.op == ("binary3", "DivNode") and mean_view.bcast_axes() == (axis,):
                msum = sum_source(imean.args[1], prod, sums)
                n_base = _through_views(imean.args[2], prod)
                if (msum is not None and msum[0] is x0 and msum[1].args[0].bcast_axes() == (axis,)
                        and not n_base.tmp and n_base.root.data.size == 1
                        and float(n_base.root.data.ravel()[0]) == float(x0.shape[axis])):
                    members += [isub, imean, msum[1], msum[2]]
                    x1, avoid = x0, True
        if not chain_closed_p(iseq, members, idiv.out_to[0]):
            continue
        for m in members:
            if m is not idiv:
                suppress(m)
        idiv.op = ("softmax", axis, avoid)
        idiv.args = [dst, x1]
    return iseq


def sum_source(t, prod, sums):
    """The (x, add-instruction, zero-fill-instruction) entry when T is, through windows only, the result of a reduce-sum."""
    while id(t) not in sums:
        p = prod.get(id(t))
        if p is None or p.node != "ViewTensorNode":
            return None
        t = p.args[1]
    return sums[id(t)]


def fuse_relu(iseq):
    """mask <- Where(x > 0 ? 1 : 0) ; out <- binary3(Mul, copy, x, mask)  ==>  out <- relu(x) when nothing else reads
    the mask and nothing is saved for backward (inference; the training graph keeps the mask for the backward)."""
    prod = producers(iseq)
    for imul in iseq:
        if imul.op != ("binary3", "MulNode"):
            continue
        dst, x, m = imul.args
        iw = prod.get(id(m))
        if (iw is None or iw.node != "Where-Operation-Node" or iw.op != ("node",) or iw.args[0] is not x
                or iw.attrs.get("scal") != 0.0 or not chain_closed_p(iseq, [iw, imul], imul.out_to[0])):
            continue
        suppress(iw)
        imul.op = ("relu",)
        imul.args = [dst, x]
    return iseq


def fuse_gemm_relu(iseq):
    """gemm_ex(a, b, out, bias) ; relu(dst, out)  ==>  gemm_ex with the activation in its epilogue, writing dst."""
    uses = use_counts(iseq)
    out, k = [], 0
    while k < len(iseq):
        i0 = iseq[k]
        j = next_launching(iseq, k)
        i1 = iseq[j] if j is not None else None
        if (i1 is not None and i0.op[0] == "gemm-ex" and i0.op[4] == 0 and not i0.op[5] and i1.op == ("relu",)
                and i1.args[1] is i0.out_to[0] and uses.get(id(i0.out_to[0])) == 1 and no_sv4bw(i0) and no_sv4bw(i1)):
            op = list(i0.op)
            op[4], op[6] = 1, i1.args[0]
            i1.op, i1.args = tuple(op), i0.args
            i1.sv4bw = []
            out += iseq[k + 1:j] + [i1]
            k = j + 1
        else:
            out.append(i0)
            k += 1
    return out


def fold_grad_adder(iseq):
    """g <- MatMul(a, b, g) ; [gp <- Permute(g) ;] grad <- Move|Add(grad, g|gp) with grad-adder-p
       ==>  one gemm writing (Move) or accumulating into (Add) the gradient itself; through a `!t` window the product
       is formed transposed: (op(a) op(b))^T = op(b)^T op(a)^T, i.e. operands swapped and both flags inverted."""
    k = 0
    while k < len(iseq):
        i0 = iseq[k]
        if i0.node == "MatMulNode" and i0.op == ("node",) and no_sv4bw(i0):
            j = k + 1
            perm = None
            if j < len(iseq) and iseq[j].node == "Permute-Node" and iseq[j].args[0] is i0.out_to[0]:
                perm, j = iseq[j], j + 1
            ad = iseq[j] if j < len(iseq) else None
            src = perm.out_to[0] if perm is not None else i0.out_to[0]
            if (ad is not None and ad.grad_adder and ad.node in ("MoveTensorNode", "AddNode") and ad.args[1] is src
                    and len(i0.args[2].shape) == 2 and not ad.args[0].bcast_axes()):
                a, b, _ = i0.args
                ta, tb = i0.attrs["ta"], i0.attrs["tb"]
                if perm is not None:
                    a, b, ta, tb = b, a, not tb, not ta
                    # an operand that was read through `!t` is now read as it lies, and the other way round
                    a, b = (untransposed(a) if not ta else a), (untransposed(b) if not tb else b)
                i0.op = ("gemm-ex", ta, tb, False, 0, ad.node == "AddNode", ad.args[0], perm is not None)
                i0.args = [a, b, i0.args[2]]
                suppress(ad)
        k += 1
    return iseq


def on_finalizing_compiling(iseq_fw, iseq_bw=None):
    """Same order as the method in fusion.lisp."""
    def rewrite(iseq):
        iseq = fuse_matmul_bias(fuse_move_arith(iseq))
        iseq, sums = fuse_zero_fill_sum(iseq)
        return fold_grad_adder(fuse_gemm_relu(fuse_relu(fuse_softmax_chain(iseq, sums))))
    iseq_fw, iseq_bw = rewrite(iseq_fw), (rewrite(iseq_bw) if iseq_bw else None)
    alias_immutable_sv4bw(iseq_fw, iseq_bw)            # last: the places are aligned with the final argument lists
    return iseq_fw, iseq_bw
