"""One wgrad+Adam grouped GEMM (config-5 layer shape, few groups) for an ncu capture: FEDB200_GEMM picks the kernel."""
import sys, numpy as np, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200.ops import gemm as gm, _native as nat
G, B, W = 4, 1024, 4096
bf = torch.bfloat16
X = torch.randn(G, B, W, device="cuda").to(bf); Wt = torch.zeros(G, W, W, device="cuda", dtype=bf)
Wm = torch.randn(G, W, W, device="cuda"); Mo = torch.zeros_like(Wm); Ve = torch.zeros_like(Wm)
groups = np.zeros(G, dtype=nat.GROUP_DTYPE); groups["opt"] = 1; groups["lr"] = 1e-3; groups["beta1"] = .9; groups["beta2"] = .999; groups["eps"] = 1e-8; groups["lr_gamma"] = 1
dg = torch.from_numpy(groups.view(np.uint8).copy()).cuda(); step = torch.ones(G, dtype=torch.int32, device="cuda")
for _ in range(3):
    gm.gemm(G, W, W, B, X, B * W, W, True, X, B * W, W, True, gm.WGRAD_OPT, W=Wm, Mom=Mo, Vel=Ve, Wsh=Wt, w_stride=W * W, ldw=W, groups=dg, step=step)
torch.cuda.synchronize(); print("ok", gm.pair_status())
