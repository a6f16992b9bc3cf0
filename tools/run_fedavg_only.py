import sys, numpy as np, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200 import FedEngine, FedProblem, prepare
d = prepare()
shards = [np.arange(c * 1000, (c + 1) * 1000, dtype=np.int32) % 8000 for c in range(int(sys.argv[1]) if len(sys.argv) > 1 else 1)]
e = FedEngine([FedProblem.custom(14, 2, shards, patience=10)], d.X_train, d.y_train, device="cuda", max_rounds=1000, precision="fp32")
for _ in range(12):
    e.enqueue_round_begin(); e.enqueue_eval_local(); e.enqueue_aggregate()
torch.cuda.synchronize()
print("ok")
