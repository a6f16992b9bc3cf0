"""Phase timeline of the fused single-launch round (fedb_round_tc_prof), warm vs after an L2 flush (the bench's timing rule).

    python tools/tc_timeline_cold.py [rows_per_client] [clients]
"""
import sys

import numpy as np
import torch

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200 import FedEngine, FedProblem, prepare  # noqa: E402
from fedb200.ops import tc  # noqa: E402

NAMES = ["start", "weights+row loaded", "X stored", "fwd1 mma", "epi1 (H1)", "fwd2 mma", "epi2a (head, H2, dz3)", "wgrad3 mma",
         "epi2b (dZ2)", "dgrad2+wgrad2 mma", "epi3 (dZ1)", "wgrad1 mma", "grad spill", "cluster sync 1", "dsmem reduce + optimizer",
         "cluster sync 2", "weights reloaded", "eval done", "tail check-in", "tail done", "fedavg + write-back", "metrics + early stop"]


def timeline(e, flush):
    prof = torch.zeros(32, dtype=torch.int64, device="cuda")
    E = 1
    for it in range(4):
        prof.zero_()
        if flush is not None:
            flush.zero_()
        tc.round_tc(e.tb, e.tc_cluster, e.X, e.y, e.idx, e.params, e.m, e.v, e.step, e.active, e.fed_state, e.loss, e.conf, e.scratch, E,
                    e.global_, e.hist_local, e.hist_global, e.tail_count, prof=prof, state_mirror=e._state_pinned)
        torch.cuda.synchronize()
    return prof.cpu().numpy()


if __name__ == "__main__":
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    clients = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    d = prepare()
    shards = [np.arange(c * rows, (c + 1) * rows, dtype=np.int32) % 8000 for c in range(clients)]
    e = FedEngine([FedProblem.custom(14, 2, shards, patience=0)], d.X_train, d.y_train, device="cuda", max_rounds=64, precision="bf16")
    e.prime()
    warm = timeline(e, None)
    cold = timeline(e, torch.empty(256 << 20, dtype=torch.uint8, device="cuda"))
    print(f"rows/client {rows}, clients {clients}, cluster {e.tc_cluster}; cycles per phase, warm vs L2-flushed")
    order = sorted((i for i in range(len(NAMES)) if warm[i]), key=lambda i: warm[i])
    pw = pc = None
    for i in order:
        if pw is not None:
            print(f"{i:2d} {NAMES[i]:28s} warm +{warm[i] - pw:7d}   cold +{cold[i] - pc:7d}   (cold - warm {cold[i] - pc - (warm[i] - pw):+6d})")
        pw, pc = warm[i], cold[i]
    print(f"total warm {warm[order[-1]] - warm[order[0]]}  cold {cold[order[-1]] - cold[order[0]]}")
