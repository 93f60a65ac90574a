#!/usr/bin/env python
"""Is the spread of the CTAs' streaming times inside the chain kernel systematic (same slow SMs every time) or noise?
Per-CTA (streamed - staged) of every gate_up / down linear of a 6-layer chain, normalised by the CTA's stage count:
correlation between layers, and the per-CTA mean.   python tools/chain_balance.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    os.environ["W4A16_GV_DEBUG"] = "1"
    import numpy as np
    import torch
    import bench
    P = bench.load_product()
    dev = torch.device("cuda", 0)
    layers = bench.build_layers(P, torch, bench.QWEN3_8B, 6, dev)
    chain = P.DecodeChain(bench.chain_steps(torch, layers, 1, dev))
    for _ in range(4):
        chain.run()
    torch.cuda.synchronize()
    a = chain.debug_stamps().astype(np.int64)       # [n, ctas, 8]
    grid = a.shape[1]
    for name, idx, nr, spr in (("gate_up", 2, 1536, 1), ("down", 3, 256, 3), ("qkv", 0, 384, 1)):
        rb = np.array([k * nr // grid for k in range(grid + 1)])
        stages = (rb[1:] - rb[:-1]) * spr
        d = np.stack([(a[4 * l + idx, :, 2] - a[4 * l + idx, :, 1]) / 1e3 for l in range(1, 6)])     # [5 layers, ctas] us
        per_stage = d / np.maximum(stages, 1)
        c = np.corrcoef(per_stage)
        m = per_stage.mean(axis=0)
        print(f"{name}: stream us mean {d.mean():.2f}  min {d.min():.2f} max {d.max():.2f}; stages/CTA {stages.min()}..{stages.max()}")
        print(f"   correlation of per-stage time between layers: mean off-diagonal {(c.sum() - 5) / 20:.2f}")
        order = np.argsort(m)
        print(f"   per-stage us: fastest CTAs {order[:6].tolist()} {np.round(m[order[:6]], 3).tolist()}  slowest {order[-6:].tolist()} {np.round(m[order[-6:]], 3).tolist()}")
        print(f"   corr(per-stage time, stages) {np.corrcoef(m, stages)[0, 1]:.2f};  finish-time spread (max - median) per layer: "
              f"{np.round(d.max(axis=1) - np.median(d, axis=1), 2).tolist()}")


if __name__ == "__main__":
    main()
