"""Config-5-shaped throughput check of the wide path on one GPU: C clients x B rows, 4 x W hidden.

    python tools/wide_bench.py [clients] [rows] [width] [rounds]
"""
import sys, time
import numpy as np, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200 import FedEngine, FedProblem, TrainHyper, synthetic_income_like
from fedb200.data.partition import dirichlet

if __name__ == "__main__":
    C = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    W = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
    R = int(sys.argv[4]) if len(sys.argv) > 4 else 5
    opt = sys.argv[5] if len(sys.argv) > 5 else "sgd"
    d = synthetic_income_like(max(20000, C * B // 4), seed=1)
    n = len(d.X_train)
    rng = np.random.default_rng(0)
    # non-IID: each client draws its B rows with a client-specific class skew
    shards = []
    pos, neg = np.flatnonzero(d.y_train == 1), np.flatnonzero(d.y_train == 0)
    for c in range(C):
        frac = rng.beta(0.5, 0.5)
        k = int(B * frac)
        shards.append(np.concatenate([rng.choice(pos, k), rng.choice(neg, B - k)]).astype(np.int32))
    hyper = TrainHyper(opt="sgd", lr=0.05, lr_step_size=0) if opt == "sgd" else TrainHyper()
    eng = FedEngine([FedProblem.custom(14, 2, shards, hidden=(W,) * 4, hyper=hyper, patience=0, init_per_client=False)],
                    d.X_train, d.y_train, d.X_test, d.y_test, device="cuda", max_rounds=R + 4)
    print("precision", eng.precision_name, "params/client", eng.problems[0].spec.n_params, "arena GB", eng.params_total * 4 / 2**30)
    eng.run(1, use_graph=False); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); h = eng.run(R, use_graph=False); e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / R
    flops = 6.0 * B * eng.problems[0].spec.n_params * C + 2.0 * B * eng.problems[0].spec.n_params * C   # train step + eval forward
    print(f"{ms:.2f} ms/round  {C * B / ms * 1e3 / 1e6:.2f} M client-samples/s  {flops / ms / 1e9:.1f} TFLOP/s (bf16, incl. eval fwd)  "
          f"acc {h['global'][0, -1, 0]:.4f}  mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GB")
