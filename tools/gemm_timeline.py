"""Phase timeline of the CTA-pair wgrad+optimizer GEMM (fedb_gemm_tma2_prof): where does a CTA's time go?"""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["FEDB200_GEMM"] = "tma2"
from fedb200.ops import gemm as gm, _native as nat
G, B, W = (int(x) for x in (sys.argv[1:4] + [16, 1024, 4096][len(sys.argv) - 1:]))
bf = torch.bfloat16
X = torch.randn(G, B, W, device="cuda").to(bf); Wt = torch.zeros(G, W, W, device="cuda", dtype=bf)
Wm = torch.randn(G, W, W, device="cuda"); Mo = torch.zeros_like(Wm); Ve = torch.zeros_like(Wm)
groups = np.zeros(G, dtype=nat.GROUP_DTYPE); groups["opt"] = 1; groups["lr"] = 1e-3; groups["beta1"] = .9; groups["beta2"] = .999; groups["eps"] = 1e-8; groups["lr_gamma"] = 1
dg = torch.from_numpy(groups.view(np.uint8).copy()).cuda(); step = torch.ones(G, dtype=torch.int32, device="cuda")
n_cta = (W // 128) * (W // 256) * G
run = lambda: gm.gemm(G, W, W, B, X, B * W, W, True, X, B * W, W, True, gm.WGRAD_OPT, W=Wm, Mom=Mo, Vel=Ve, Wsh=Wt, w_stride=W * W, ldw=W, groups=dg, step=step)
for _ in range(2): run()
gm.PROF = torch.zeros(n_cta * 8, dtype=torch.int64, device="cuda")
run(); torch.cuda.synchronize()
t = gm.PROF.cpu().numpy().reshape(n_cta, 8).astype(np.float64)
gm.PROF = None
t0 = t[:, 0].min()
names = ["main loop (start -> accumulators done)", "stage pass 0", "rows pass 0", "stage pass 1", "rows pass 1"]
print(f"kernel span {(t[:, 5].max() - t0) / 1e3:.1f} us, {n_cta} CTAs, {len(np.unique(t[:, 6]))} SMs")
for k, n in enumerate(names):
    d = (t[:, k + 1] - t[:, k]) / 1e3
    print(f"{n:42s} mean {d.mean():7.2f} us  p10 {np.percentile(d, 10):7.2f}  p50 {np.percentile(d, 50):7.2f}  p90 {np.percentile(d, 90):7.2f}")
d = (t[:, 5] - t[:, 0]) / 1e3
print(f"{'CTA lifetime':42s} mean {d.mean():7.2f} us  p10 {np.percentile(d, 10):7.2f}  p50 {np.percentile(d, 50):7.2f}  p90 {np.percentile(d, 90):7.2f}")
# per-SM concurrency: time-average number of resident CTAs and of CTAs in a rows phase
span = t[:, 5].max() - t0
print(f"mean resident CTAs per SM {((t[:, 5] - t[:, 0]).sum() / span / len(np.unique(t[:, 6]))):.2f};"
      f" mean CTAs streaming rows per SM {(((t[:, 3] - t[:, 2]) + (t[:, 5] - t[:, 4])).sum() / span / len(np.unique(t[:, 6]))):.2f}")
