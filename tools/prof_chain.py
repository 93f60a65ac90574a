#!/usr/bin/env python
"""ncu target for the persistent chain kernel: `--layers` Qwen3-8B layers (qkv, o, gate_up, down with the model's data
flow, bench.chain_steps) as one DecodeChain, run `--reps` times with plain launches.
    ncu --set full --import-source on -k regex:chain_kernel -s 1 -c 1 ... python tools/prof_chain.py"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=1)
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    import torch
    import bench
    P = bench.load_product()
    dev = torch.device("cuda", 0)
    layers = bench.build_layers(P, torch, bench.QWEN3_8B, args.layers, dev)
    chain = P.DecodeChain(bench.chain_steps(torch, layers, args.m, dev))
    print("kernel:", chain.kernel())
    for _ in range(args.reps):
        chain.run()
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
