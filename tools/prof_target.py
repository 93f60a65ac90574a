#!/usr/bin/env python
"""Small, fixed launch sequence for ncu: the four Qwen3-8B projection shapes at the given M,
`--reps` passes over `--copies` distinct weight copies per shape (no CUDA graph, so every
kernel is a plain launch ncu can attribute).  Usage under gpurun:
    python tools/prof_target.py --m 1 > gpurun_out/plain.log 2>&1 &&
    ncu --set full ... python tools/prof_target.py --m 1
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

SHAPES = [("qkv", 4096, 6144), ("o", 4096, 4096), ("gate_up", 4096, 24576), ("down", 12288, 4096)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=1)
    ap.add_argument("--copies", type=int, default=3)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--shapes", nargs="*", default=None)
    args = ap.parse_args()
    import torch
    import conftest
    P = conftest._load_product()
    dev = torch.device("cuda", 0)
    cfg = P.AwqConfig.int4(128)
    lins = []
    for name, K, N in SHAPES:
        if args.shapes and name not in args.shapes:
            continue
        for c in range(args.copies):
            qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev)
            qzeros = torch.randint(-2**31, 2**31 - 1, (K // 128, N // 8), dtype=torch.int32, device=dev)
            scales = (torch.rand((K // 128, N), device=dev) * 0.018 + 0.002).to(torch.float16)
            lin = cfg.create_linear(K, N, False, dev)
            lin.load_weights({"qweight": qweight, "qzeros": qzeros, "scales": scales})
            lins.append((name, K, N, lin))
    torch.cuda.synchronize()
    M = args.m
    bufs = {}
    for name, K, N, lin in lins:
        bufs.setdefault((K, N), (torch.randn((M, K), device=dev).to(torch.bfloat16),
                                 torch.empty((M, N), device=dev, dtype=torch.bfloat16)))
    for _ in range(args.reps):
        for name, K, N, lin in lins:
            x, y = bufs[(K, N)]
            lin.forward_into(x, y)
    torch.cuda.synchronize()
    print("ok", P.kernel_launches(), "launches; kernel", lins[0][3].kernel_for(M))


if __name__ == "__main__":
    main()
