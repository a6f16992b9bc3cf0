"""Phase timeline of one mini-batch step of k_fit_tcx (globaltimer ns, thread 0 of each of the client's four CTAs) and the
per-epoch cost of a whole fit.

    python tools/fitx_timeline.py [hidden1 hidden2]
"""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200 import FedEngine, FedProblem, prepare  # noqa: E402
from fedb200.data.partition import contiguous  # noqa: E402
from fedb200.ops import kernels as K  # noqa: E402

NAMES = ["batch begins", "X stored", "fwd1 mma", "epi1 (H1)", "fwd2 mma", "epi2a (H2, partial logits)", "logits exchanged (barrier)",
         "loss, dz3", "wgrad3 mma", "epi2b (dZ2)", "dgrad2+wgrad2 mma", "epi3 (dZ1)", "wgrad1 mma", "all tiles done", "layer-1 partials sent",
         "slice optimizer done", "cluster wait", "layer-1 optimizer done",
         "  slice: cluster arrive done", "  slice: W2 + b2 + W3 updated (state in TMEM)", "  slice: stores complete"]

if __name__ == "__main__":
    hidden = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (50, 400)
    d = prepare(with_mean=False)
    e = FedEngine([FedProblem.sklearn_like(d.n_features, contiguous(len(d.X_train), 4), hidden=hidden)], d.X_train, d.y_train, device="cuda",
                  max_rounds=2)
    h = e.problems[0].hyper
    prof = torch.zeros(4 * 32, dtype=torch.int64, device="cuda")
    for it in range(3):
        e.reset_fit_state()
        prof.zero_()
        act = torch.ones_like(e.active)
        K.fit_tcx(e.tb, e.params, e.m, e.v, e.X, e.y, e.idx, e.idx_epoch, e.step, act, e.fed_state, e.epoch, e.no_improve, e.best_loss,
                  e.last_loss, e.loss, h.tol, h.n_iter_no_change, h.max_iter, 2, True, e.seed + 0x5EED, prof)
        torch.cuda.synchronize()
    p = prof.cpu().numpy().reshape(4, 32)
    t0 = p[:, 0].min()
    print(f"model 14-{hidden[0]}-{hidden[1]}-1, mini-batch 200 (two 128-row tiles; stamps 1..12 are the first tile); ns since the batch began")
    for i, nme in enumerate(NAMES):
        print(f"{i:2d} {nme:32s}" + "".join(f"{int(v - t0):>8d}" for v in p[:, i]))
    e.reset_fit_state()
    e.init_params()
    torch.cuda.synchronize()
    t = time.perf_counter()
    n = e.fit_local()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    print(f"whole fit: epochs {n.tolist()}, {dt * 1e3:.1f} ms = {dt * 1e6 / n.max():.0f} us per epoch of {e.n_batches} mini-batches")
