"""TFLOP/s of the grouped tcgen05 GEMM per mode (config-5 layer shape by default)."""
import sys
import numpy as np, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200.ops import gemm as gm, _native as nat

def timed(fn, n=10):
    fn(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n

if __name__ == "__main__":
    G, B, W = (int(x) for x in (sys.argv[1:4] + [16, 1024, 4096][len(sys.argv) - 1:]))
    bf = torch.bfloat16
    X = torch.randn(G, B, W, device="cuda").to(bf); Wt = (torch.randn(G, W, W, device="cuda") / 64).to(bf)
    Y = torch.empty(G, B, W, device="cuda", dtype=bf); bias = torch.zeros(G, W, device="cuda")
    fl = 2.0 * G * B * W * W
    t = timed(lambda: gm.gemm(G, B, W, W, X, B * W, W, False, Wt, W * W, W, False, gm.FWD_BIAS_RELU, C=Y, c_stride=B * W, ldc=W, bias=bias, bias_stride=W))
    print(f"fwd   {t:.3f} ms  {fl / t / 1e9:.0f} TFLOP/s")
    t = timed(lambda: gm.gemm(G, B, W, W, X, B * W, W, False, Wt, W * W, W, True, gm.DGRAD_MASK, C=Y, c_stride=B * W, ldc=W, act=X, act_stride=B * W, ldact=W))
    print(f"dgrad {t:.3f} ms  {fl / t / 1e9:.0f} TFLOP/s")
    Wm = torch.randn(G, W, W, device="cuda"); Mo = torch.zeros_like(Wm); Ve = torch.zeros_like(Wm)
    groups = np.zeros(G, dtype=nat.GROUP_DTYPE); groups["opt"] = 1; groups["lr"] = 1e-3; groups["beta1"] = .9; groups["beta2"] = .999; groups["eps"] = 1e-8; groups["lr_gamma"] = 1
    dg = torch.from_numpy(groups.view(np.uint8).copy()).cuda(); step = torch.ones(G, dtype=torch.int32, device="cuda")
    t = timed(lambda: gm.gemm(G, W, W, B, X, B * W, W, True, X, B * W, W, True, gm.WGRAD_OPT, W=Wm, Mom=Mo, Vel=Ve, Wsh=Wt, w_stride=W * W, ldw=W, groups=dg, step=step))
    print(f"wgrad+adam {t:.3f} ms  {fl / t / 1e9:.0f} TFLOP/s   (epilogue traffic {G * W * W * 26 / t / 1e6:.0f} GB/s)")
    a, b = X[0].float(), Wt[0].float()
    t = timed(lambda: torch.matmul(X, Wt.transpose(1, 2)))
    print(f"torch bmm (cuBLAS) {t:.3f} ms  {fl / t / 1e9:.0f} TFLOP/s")
