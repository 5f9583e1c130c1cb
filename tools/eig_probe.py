import sys, time, torch
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
torch.manual_seed(0)
dev = torch.device("cuda")
# covariance-like matrix: low-rank signal + noise
F = 4000
X = torch.randn(n, 40, device=dev) @ (torch.randn(40, F, device=dev) * torch.logspace(0, -2, 40, device=dev)[:, None]) + 0.05 * torch.randn(n, F, device=dev)
X -= X.mean(1, keepdim=True)
A = (X @ X.T / (F - 1)).float()
del X
torch.cuda.synchronize()
for dt in (torch.float64,):
    t = time.time(); w, v = torch.linalg.eigh(A.to(dt)); torch.cuda.synchronize(); print("eigh", dt, n, "%.2f s" % (time.time() - t), w[-3:].tolist())
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conan_b200 import stage3
for k in (3, 10):
    torch.cuda.synchronize(); t = time.time(); wk, vk = stage3.top_eigenpairs(A, k); torch.cuda.synchronize()
    print("top_eigenpairs k=%d %.3f s" % (k, time.time() - t), wk[:3].tolist(), "max rel diff to eigh", ((wk - w.flip(0)[:k]).abs() / w[-1]).max().item())
