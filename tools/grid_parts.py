import sys, time, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200 import prepare
from fedb200.fed.tune import run_grid, HIDDEN_GRID, LR_GRID
d = prepare(with_mean=False)
sets = {"all": list(HIDDEN_GRID), "heavy (400,200),(200,400)": [(400, 200), (200, 400)], "(100,400)": [(100, 400)], "(100,50)": [(100, 50)],
        "(50,),(100,)": [(50,), (100,)], "tc shapes": [(50, 50), (50, 100), (50, 200), (50, 400)],
        "all but heavy": [h for h in HIDDEN_GRID if h not in ((400, 200), (200, 400))]}
sel = sys.argv[1:]
for name, hg in sets.items():
    if sel and not any(x in name for x in sel):
        continue
    for it in range(2):
        tm = {}
        torch.cuda.synchronize(); t = time.perf_counter()
        out = run_grid(d, 4, hidden_grid=hg, device="cuda", timing=tm)
        torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"{name:28s} wall {dt*1e3:8.1f} ms  device {tm.get('device_ms', 0):8.1f} ms  epochs(SIMT) {tm.get('epochs')}  tc clients {tm.get('tc_clients')}/{tm.get('clients')}")
