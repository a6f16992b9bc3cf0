"""Per-CTA phase timeline of k_round_tcx (globaltimer ns; every CTA of the first cluster stamps its thread 0).

    python tools/tcx_timeline.py [rows_per_client] [clients] [cold]      ("cold": flush L2 with a 256 MiB memset before the launch)
"""
import sys

import numpy as np
import torch

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200 import FedEngine, FedProblem, prepare  # noqa: E402
from fedb200.ops import tc  # noqa: E402

NAMES = {0: "start", 1: "weights converted", 2: "X stored", 3: "fwd1 mma", 4: "epi1", 5: "fwd2 mma", 6: "epi2a", 7: "wgrad3 mma", 8: "epi2b",
         9: "dgrad2+wgrad2 mma", 10: "epi3", 11: "wgrad1 mma", 12: "grads staged", 13: "partials in", 14: "optimizer done", 15: "barrier",
         16: "weights converted", 17: "eval done", 18: "tail begins", 20: "write-back done", 21: "metrics done", 19: "end",
         22: "metrics: entry", 23: "metrics: per-client done", 24: "metrics: reduced", 25: "metrics: state written"}
ORDER = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 20, 22, 23, 24, 25, 21, 19]

if __name__ == "__main__":
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    clients = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    d = prepare()
    shards = [np.arange(c * rows, (c + 1) * rows, dtype=np.int32) % 8000 for c in range(clients)]
    e = FedEngine([FedProblem.custom(14, 2, shards, patience=0)], d.X_train, d.y_train, device="cuda", max_rounds=64, precision="tc32")
    e.prime()
    prof = torch.zeros(16 * 32, dtype=torch.int64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if (len(sys.argv) > 3 and sys.argv[3] == "cold") else None
    for it in range(4):
        prof.zero_()
        if flush is not None:
            flush.zero_()
        tc.round_tc(e.tb, e.tc_cluster, e.X, e.y, e.idx, e.params, e.m, e.v, e.step, e.active, e.fed_state, e.loss, e.conf, e.scratch,
                    1, e.global_, e.hist_local, e.hist_global, e.tail_count, prof, variant="x2")
        torch.cuda.synchronize()
    p = prof.cpu().numpy().reshape(16, 32)
    cs = e.tc_cluster
    t0 = p[:cs, 0].min()
    print(f"rows/client {rows}, clients {clients}, cluster {cs}; ns since the first CTA's start (4th launch)")
    print("stamp".ljust(22) + "".join(f"{r:>6d}" for r in range(cs)))
    for i in ORDER:
        if p[:cs, i].any():
            print(NAMES[i].ljust(22) + "".join(f"{(int(v - t0) if v else -1):>6d}" for v in p[:cs, i]))
