"""Event-timed comparison of round variants on one GPU + the fused kernel's phase timeline."""
import sys

import numpy as np
import torch

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200 import FedEngine, FedProblem, prepare  # noqa: E402
from fedb200.ops import kernels as K, tc  # noqa: E402


def timed(fn, n=300, graph=True):
    fn(); torch.cuda.synchronize()
    if graph:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        run = g.replay
    else:
        run = fn
    for _ in range(20):
        run()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record()
    for _ in range(n):
        run()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3


if __name__ == "__main__":
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    clients = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    PREC = sys.argv[3] if len(sys.argv) > 3 else "tc32"
    d = prepare()
    shards = [np.arange(c * rows, (c + 1) * rows, dtype=np.int32) % 8000 for c in range(clients)]
    e = FedEngine([FedProblem.custom(14, 2, shards, patience=0)], d.X_train, d.y_train, device="cuda", max_rounds=100000, precision=PREC)
    e.prime()
    body = lambda ev: tc.local_round_tc(e.tb, e.tc_cluster, e.X, e.y, e.idx, e.params, e.m, e.v, e.step, e.active, e.fed_state, e.loss,
                                        e.conf, e.scratch, 0, 1, ev, variant=e.tc_variant)
    print("empty-ish kernel (round_begin)      us:", round(timed(e.enqueue_round_begin), 2))
    print("tc body, train only                 us:", round(timed(lambda: body(False)), 2))
    print("tc body, train + eval               us:", round(timed(lambda: body(True)), 2))
    print("fedavg kernel alone (begin_next)    us:", round(timed(lambda: e.enqueue_aggregate()), 2))
    print("body + fedavg (2 launches)          us:", round(timed(lambda: (body(True), e.enqueue_aggregate())), 2))
    print("fused round (1 launch)              us:", round(timed(e.enqueue_round), 2))
    print("fused round, eager launches         us:", round(timed(e.enqueue_round, graph=False), 2))
    prof = torch.zeros(32, dtype=torch.int64, device="cuda")
    for _ in range(3):
        prof.zero_()
        tc.round_tc(e.tb, e.tc_cluster, e.X, e.y, e.idx, e.params, e.m, e.v, e.step, e.active, e.fed_state, e.loss, e.conf, e.scratch,
                    1, e.global_, e.hist_local, e.hist_global, e.tail_count, prof, variant=e.tc_variant)
        torch.cuda.synchronize()
    p = prof.cpu().numpy()
    print("timeline (cycles from start):", {i: int(p[i] - p[0]) for i in range(24) if p[i]})
