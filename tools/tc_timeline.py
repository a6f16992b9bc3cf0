"""Phase timeline of k_local_round_tc from in-kernel clock64 stamps (CTA 0, thread 0).

    python tools/tc_timeline.py [rows_per_client] [clients] [tc32|bf16]
"""
import sys

import numpy as np
import torch

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200 import FedEngine, FedProblem, prepare  # noqa: E402
from fedb200.ops import tc  # noqa: E402

NAMES = ["start", "weights+row loaded", "X stored", "fwd1 mma", "epi1 (H1)", "fwd2 mma", "epi2a (head, H2, dz3)", "wgrad3 mma",
         "epi2b (dZ2)", "dgrad2+wgrad2 mma", "epi3 (dZ1)", "wgrad1 mma", "grad spill", "cluster sync 1", "dsmem reduce + optimizer",
         "cluster sync 2", "weights reloaded", "eval done"]

if __name__ == "__main__":
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    clients = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    PREC = sys.argv[3] if len(sys.argv) > 3 else "tc32"
    d = prepare()
    shards = [np.arange(c * rows, (c + 1) * rows, dtype=np.int32) % 8000 for c in range(clients)]
    e = FedEngine([FedProblem.custom(14, 2, shards, patience=0)], d.X_train, d.y_train, device="cuda", max_rounds=8, precision=PREC)
    prof = torch.zeros(32, dtype=torch.int64, device="cuda")
    for it in range(3):
        prof.zero_()
        tc.local_round_tc(e.tb, e.tc_cluster, e.X, e.y, e.idx, e.params, e.m, e.v, e.step, e.active, e.fed_state, e.loss, e.conf,
                          e.scratch, 0, 1, True, prof, variant=e.tc_variant)
        torch.cuda.synchronize()
    p = prof.cpu().numpy()
    t0 = p[0]
    print(f"rows/client {rows}, clients {clients}, cluster {e.tc_cluster}; cycles since start (last launch, warm)")
    prev = t0
    for i, n in enumerate(NAMES):
        if p[i]:
            print(f"{i:2d} {n:28s} +{p[i] - prev:7d}  @{p[i] - t0:8d}")
            prev = p[i]
