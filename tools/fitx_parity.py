"""How far the tensor-core fit kernel (k_fit_tcx) is from the fp32 CUDA-core kernels after a few epochs, layer by layer, for
plain SGD (errors propagate linearly: anything above ~1e-6 is a defect) and for sklearn's Adam (whose first steps are
+-lr * g / |g| and amplify rounding-level differences of small gradients).

    python tools/fitx_parity.py [epochs] [hidden1 hidden2]
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200 import FedEngine, FedProblem, prepare  # noqa: E402
from fedb200.data.partition import contiguous  # noqa: E402
from fedb200.fed.problem import TrainHyper  # noqa: E402


def pair(d, hidden, hyper, rows=4000, clients=2):
    def mk():
        return FedEngine([FedProblem.sklearn_like(d.n_features, contiguous(rows, clients), hidden=hidden, hyper=hyper)], d.X_train[:rows],
                         d.y_train[:rows], device="cuda", max_rounds=2)
    os.environ["FEDB200_FIT"] = "launches"
    simt = mk()
    del os.environ["FEDB200_FIT"]
    tcx = mk()
    assert simt._fit_tc_mask() is None and tcx._fit_tc_mask() is not None and bool(tcx._fit_tc_mask().all())
    return simt, tcx


if __name__ == "__main__":
    epochs = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    hidden = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (50, 400)
    d = prepare(with_mean=False)
    for name, hyper in (("sgd", TrainHyper(opt="sgd", lr=0.05, alpha=1e-4, lr_step_size=0, lr_gamma=1.0, batch_size=200, max_iter=300)),
                        ("sgd+momentum", TrainHyper(opt="sgd", lr=0.02, momentum=0.9, alpha=1e-4, lr_step_size=0, lr_gamma=1.0, batch_size=200,
                                                    max_iter=300)),
                        ("adam_sklearn", TrainHyper(opt="adam_sklearn", lr=0.004, alpha=1e-4, lr_step_size=0, lr_gamma=1.0, batch_size=200,
                                                    max_iter=300))):
        for shuffle in (False, True):
            simt, tcx = pair(d, hidden, hyper)
            for e in (simt, tcx):
                e.reset_fit_state()
                e.enqueue_round_begin()
                for _ in range(epochs):
                    e.active.fill_(1)
                    e.fit_local(max_iter=1, shuffle=shuffle)
            torch.cuda.synchronize()
            l0, l1 = simt.last_loss.cpu().numpy(), tcx.last_loss.cpu().numpy()
            print(f"{name:13s} shuffle={int(shuffle)} epochs={epochs}: loss {l0.tolist()} rel diff {(np.abs(l1 - l0) / np.abs(l0)).tolist()}")
            spec = simt.problems[0].spec
            P = spec.slab_size
            for k in range(2):
                a, b = simt.params[k * P:(k + 1) * P], tcx.params[k * P:(k + 1) * P]
                parts = []
                for l in range(spec.n_layers):
                    wa, wb = spec.weight_view(a, l), spec.weight_view(b, l)
                    ba, bb = spec.bias_view(a, l), spec.bias_view(b, l)
                    parts.append(f"W{l + 1} {float((wa - wb).abs().max()):.2e}/{float(wa.abs().max()):.2e} b{l + 1} {float((ba - bb).abs().max()):.2e}/{float(ba.abs().max()):.2e}")
                dd = (a - b).abs() / a.abs().max()
                print(f"    client {k}: max |diff| / max |value|: " + "  ".join(parts))
                print(f"              relative to the largest weight: mean {float(dd.mean()):.2e}  q99 {float(dd.quantile(0.99)):.2e}  q999 {float(dd.quantile(0.999)):.2e}"
                      f"  m: {float((simt.m[k * P:(k + 1) * P] - tcx.m[k * P:(k + 1) * P]).abs().max() / simt.m[k * P:(k + 1) * P].abs().max()):.2e}")
