#!/usr/bin/env python3
"""Times conan_gemm_3xtf32 on covariance-sized operands (symmetric 16 896 x 16 896, K = 50 048)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conan_b200 import stage3
M = int(sys.argv[1]) if len(sys.argv) > 1 else 16896
K = int(sys.argv[2]) if len(sys.argv) > 2 else 50048
dev = torch.device("cuda")
x = torch.randn((M, K), device=dev, dtype=torch.float32) * 100.0
hi = x.to(torch.float16)
lo = (x - hi.float()).to(torch.float16)
del x
out = torch.empty((M, M), device=dev, dtype=torch.float32)
for sym in (True, False):
    stage3.gemm_3xtf32(hi, lo, hi, lo, K, symmetric=sym, out=out)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(2):
        stage3.gemm_3xtf32(hi, lo, hi, lo, K, symmetric=sym, out=out)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 2
    nt = (M + 127) // 128
    tiles = nt * (nt + 1) / 2 if sym else nt * nt
    mma = 3 * 2.0 * tiles * 128 * 128 * K
    print("symmetric=%s %.2f ms  %.0f f16 TFLOP/s issued (%.0f%% of 1664)" % (sym, ms, mma / ms / 1e9, mma / ms / 1e9 / 16.64))
