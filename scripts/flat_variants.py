"""Time `sum :axis t` (reduce_flat_bulk_kernel) and the two other C2 reductions the way bench.py times a case (20 launches
over rotating buffer sets in one CUDA graph, best of 3) under the launch shape given by WF2CU_FLAT_* (read once per process)."""
import ctypes as C, importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
wf = importlib.import_module("cl-waffe2_b200"); wf.init(0)
lib = wf._lib.load(); chk = wf._lib.check
sw = bench.GpuSweep(wf, 4096)
def timed(name, reps=20):
    def many():
        for _ in range(reps):
            sw.cursor = (sw.cursor + 1) % bench.NSETS
            sw.run_case(name, sw.sets[sw.cursor])
    many()
    ex = C.c_void_p(); chk(lib.wf2cu_graph_begin()); many(); chk(lib.wf2cu_graph_end(C.byref(ex)))
    chk(lib.wf2cu_graph_launch(ex)); best = None
    for _ in range(5):
        chk(lib.wf2cu_timer_start()); chk(lib.wf2cu_graph_launch(ex)); t = C.c_float(); chk(lib.wf2cu_timer_stop(C.byref(t)))
        best = t.value if best is None else min(best, t.value)
    chk(lib.wf2cu_graph_destroy(ex))
    return best / reps * 1e3
tag = " ".join(f"{k[11:]}={os.environ[k]}" for k in sorted(os.environ) if k.startswith("WF2CU_FLAT_")) or "defaults"
print(f"{tag:32s} sum_all {timed('sum_all'):6.2f} us   sum_axis1 {timed('sum_axis1'):6.2f}   sum_axis0 {timed('sum_axis0'):6.2f}", flush=True)
