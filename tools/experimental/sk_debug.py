#!/usr/bin/env python
"""Which k blocks does the skinny kernel get wrong?  x = one-hot rows (token m has a 1 at k = 64 m + 3)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, conftest
P = conftest._load_product()
P.set_dispatch("skinny")
dev = torch.device("cuda", 0)
K, N, g = int(sys.argv[1]), int(sys.argv[2]), 128
cfg = P.AwqConfig.int4(g)
qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev)
qzeros = torch.randint(-2**31, 2**31 - 1, (K // g, N // 8), dtype=torch.int32, device=dev)
scales = (torch.rand((K // g, N), device=dev) * 0.018 + 0.002).to(torch.float16)
lin = cfg.create_linear(K, N, False, dev)
lin.load_weights({"qweight": qweight, "qzeros": qzeros, "scales": scales})
W = lin.dequantize().float()
M = K // 64
x = torch.zeros((M, K), device=dev, dtype=torch.bfloat16)
for m in range(M):
    x[m, 64 * m + 3] = 1.0
y = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16)
lin.forward_into(x, y)
torch.cuda.synchronize()
ref = x.float() @ W
for m in range(M):
    err = (y[m].float() - ref[m]).abs()
    bad = (err > 1e-3).nonzero().flatten().tolist()
    print(f"kblock {m}: max_err={err.max().item():.4g} bad_cols={len(bad)} first={bad[:8]}")
