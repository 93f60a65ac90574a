#!/usr/bin/env python
"""Per-linear timeline of the persistent chain kernel (W4A16_GV_DEBUG stamps of w4a16_chain.cu) on the Qwen3-8B decode
step: python tools/chain_timeline.py [--m 1] [--layers 6].  Prints the last layer's four linears: min / median / max over
CTAs of {linear started, activations staged, weights streamed, y written + barrier passed}, relative to the start of qkv."""
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
    ap.add_argument("--layers", type=int, default=6)
    args = ap.parse_args()
    os.environ["W4A16_GV_DEBUG"] = "1"
    import numpy as np
    import torch
    import conftest
    P = conftest._load_product()
    dev = torch.device("cuda", 0)
    cfg = P.AwqConfig.int4(128)
    M = args.m
    import bench
    layers, names = [], []
    for c in range(args.layers):
        for name, K, N in SHAPES:
            qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev)
            qzeros = torch.randint(-2**31, 2**31 - 1, (K // 128, N // 8), dtype=torch.int32, device=dev)
            scales = (torch.rand((K // 128, N), device=dev) * 0.018 + 0.002).to(torch.float16)
            lin = cfg.create_linear(K, N, False, dev)
            lin.load_weights({"qweight": qweight, "qzeros": qzeros, "scales": scales})
            layers.append((name, K, N, K, N, None, lin))
            names.append(name)
    steps = bench.chain_steps(torch, layers, M, dev)
    chain = P.DecodeChain(steps)
    print("kernel:", chain.kernel())
    st = torch.cuda.Stream(dev)
    with torch.cuda.stream(st):
        for _ in range(3):
            chain.run()
        st.synchronize()
    a = chain.debug_stamps().astype(np.int64)       # [n, ctas, 4]
    n = len(steps)
    t0 = a[n - 4, :, 0].min()
    labels = [(0, "started"), (4, "hint-seen"), (5, "x-valid"), (7, "x-all"), (1, "staged"), (2, "streamed"), (3, "written")]
    for o in range(n - 4, n):
        print(f"-- {names[o]}")
        for i, nm in labels:
            if a[o, :, i].max() == 0:
                continue
            v = (a[o, :, i] - t0) / 1e3
            print(f"   {nm:9s} min {v.min():8.2f}  med {np.median(v):8.2f}  max {v.max():8.2f} us")
        print(f"   re-reads of the first pieces: med {np.median(a[o, :, 6]):.0f} max {a[o, :, 6].max()}")
    total = (a[n - 1, :, 3].max() - a[0, :, 0].min()) / 1e3
    print(f"whole chain ({n} linears): {total:.2f} us = {total / args.layers:.2f} us per layer")


if __name__ == "__main__":
    main()
