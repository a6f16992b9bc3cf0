"""A few epochs of the grid's two widest shapes on the CUDA-core kernels (for ncu: -k regex:k_fwd_bwd)."""
import sys
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200 import prepare
from fedb200.fed.tune import run_grid
d = prepare(with_mean=False)
out = run_grid(d, 4, hidden_grid=[(400, 200), (200, 400)], max_iter_grid=(4,), device="cuda")
print(len(out["results"]), out["best_metrics"]["accuracy"])
