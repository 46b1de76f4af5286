"""What the exchange kernel itself costs: the C4 MLP step (batch 1024, Adam) on ONE GPU with (a) the plain Adam launch and
(b) the peer-memory exchange kernel attached to a world of one rank (nothing on the wire).  WF2CU_P2P_LL=0 for the pull kernel."""
import ctypes as C, importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
wf = importlib.import_module("cl-waffe2_b200"); wf.init(0)
nn = importlib.import_module("cl-waffe2_b200.nn"); mb = importlib.import_module("cl-waffe2_b200.mlp_bench")
lib = wf._lib.load(); chk = wf._lib.check
X, Y = mb.shard_data(0, 1024)
def run(p2p, steps=300):
    model = nn.MLPSequence(mb.DIMS, "relu", seed=1234)
    step = nn.CompiledStep(model, "softmax-cross-entropy-sum", nn.Adam(model.arena, lr=1e-3), X.shape, Y.shape, world=1, exchange="p2p" if p2p else "nccl")
    if p2p:
        step.p2p_single = True
        nn.p2p_attach(model.arena, 0, 1, lambda b: [b])
    step.set_inputs(X, Y)
    for _ in range(30): step.step()
    chk(lib.wf2cu_sync())
    best = None
    for _ in range(3):
        chk(lib.wf2cu_timer_start())
        for _ in range(steps): step.step()
        ms = C.c_float(); chk(lib.wf2cu_timer_stop(C.byref(ms)))
        best = ms.value if best is None else min(best, ms.value)
    loss = step.loss_value(); step.close()
    if p2p: nn.p2p_detach()
    return best / steps * 1e3, loss
a, la = run(False); b, lb = run(True)
print(f"plain Adam step {a:.2f} us (loss {la:.6f})   exchange kernel on a world of one {b:.2f} us (loss {lb:.6f})   LL={os.environ.get('WF2CU_P2P_LL', '1')}")
