"""Where the end-to-end step time goes: graph variants with/without the host copies."""
import sys, time
import numpy as np, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200 import prepare
from fedb200.parallel.dist_engine import make_local_engine

d = prepare()
e = make_local_engine(d, [np.arange(1000)], max_rounds=200000)
e.prime()
xh, yh = e.pinned_train_inputs()
sp = torch.empty(e.fed_state.numel(), dtype=torch.uint8).pin_memory()

def variant(h2d_x, h2d_y, d2h):
    def body():
        if h2d_x: e.X[: e.n_train].copy_(xh, non_blocking=True)
        if h2d_y: e.y[: e.n_train].copy_(yh, non_blocking=True)
        e.enqueue_round()
        if d2h: sp.copy_(e.fed_state, non_blocking=True)
    body(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        body()
    for _ in range(50):
        g.replay(); torch.cuda.current_stream().synchronize()
    s, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 500
    torch.cuda.synchronize(); t0 = time.perf_counter(); s.record()
    for _ in range(n):
        g.replay(); torch.cuda.current_stream().synchronize()
    en.record(); torch.cuda.synchronize()
    return s.elapsed_time(en) / n * 1e3, (time.perf_counter() - t0) / n * 1e6

for name, v in (("kernel + sync", (0, 0, 0)), ("+ H2D X (64 KB)", (1, 0, 0)), ("+ H2D X + y", (1, 1, 0)), ("+ H2D X + y + D2H", (1, 1, 1)), ("D2H only", (0, 0, 1))):
    ev, wall = variant(*v)
    print(f"{name:24s} device-event {ev:6.1f} us   wall {wall:6.1f} us")
