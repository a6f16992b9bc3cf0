"""The 90-point grid on this machine against the reference's table (tests/golden.py = SURVEY.md 6.2).

    python tools/grid_parity.py [feistel|mt19937] [clients]
"""
import sys
import time

import numpy as np

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
sys.path.insert(0, __import__("os").path.join(__import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))), "tests"))
import golden  # noqa: E402
from fedb200 import prepare  # noqa: E402
from fedb200.fed.tune import HIDDEN_GRID, LR_GRID, accuracy_table, run_grid  # noqa: E402

if __name__ == "__main__":
    stream = sys.argv[1] if len(sys.argv) > 1 else "feistel"
    clients = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    d = prepare(with_mean=False)
    t0 = time.perf_counter()
    out = run_grid(d, clients, shuffle_stream=stream)
    dt = time.perf_counter() - t0
    t = accuracy_table(out["results"], HIDDEN_GRID, LR_GRID)
    diff = np.abs(t - golden.GRID_ACCURACY)
    np.set_printoptions(precision=3, suppress=True, linewidth=160)
    print(f"shuffle stream {stream}, {clients} clients, {dt:.2f} s; best {out['best_params']} {out['best_metrics']['accuracy']:.4f}")
    print(t)
    low = diff[:, :7]
    print(f"|ours - reference|: lr <= 0.05 columns: max {low.max():.3f} mean {low.mean():.4f}, cells > 0.01: {(low > 0.01).sum()} of {low.size}, "
          f"> 0.02: {(low > 0.02).sum()}; all columns: max {diff.max():.3f} mean {diff.mean():.4f}")
