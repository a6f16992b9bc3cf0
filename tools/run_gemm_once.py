import sys, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from fedb200.ops import gemm as gm
G, B, W = 4, 1024, 4096
bf = torch.bfloat16
X = torch.randn(G, B, W, device="cuda").to(bf); Wt = (torch.randn(G, W, W, device="cuda") / 64).to(bf)
Y = torch.empty(G, B, W, device="cuda", dtype=bf); bias = torch.zeros(G, W, device="cuda")
for _ in range(4):
    gm.gemm(G, B, W, W, X, B * W, W, False, Wt, W * W, W, False, gm.FWD_BIAS_RELU, C=Y, c_stride=B * W, ldc=W, bias=bias, bias_stride=W)
torch.cuda.synchronize(); print("ok")
