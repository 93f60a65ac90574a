#!/usr/bin/env python
"""A/B of launch-time switches on the headline decode step (Qwen3-8B, 36 layers x 4 linears, M tokens): the layers are
built once; for every configuration the environment is set, the step is captured into a fresh CUDA graph (the switches
are read at launch, i.e. at capture) and timed.  A development probe, not a test.  Tuning knobs are compiled out of the
shipped library: build it with `make -C vllm-rust_b200/csrc EXTRA_NVFLAGS=-DW4A16_EXPERIMENT_ENV` first.
    python tools/decode_ab.py --cfg "W4A16_L2PF=0" --cfg "W4A16_L2PF=1" --cfg "W4A16_L2PF=2,W4A16_L2PF_HOPS=2"
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", action="append", default=[])
    ap.add_argument("--m", type=int, default=1)
    ap.add_argument("--layers", type=int, default=36)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--chain", action="store_true", help="every linear reads what the previous one wrote (M = 1)")
    ap.add_argument("--api", default="linear", choices=["linear", "chain"], help="chain: one DecodeChain (persistent launch) per step")
    args = ap.parse_args()
    import torch
    P = bench.load_product()
    dev = torch.device("cuda", 0)
    layers = bench.build_layers(P, torch, bench.QWEN3_8B, args.layers, dev)
    M = args.m
    xs = {K: torch.randn((M, K), device=dev).to(torch.bfloat16) for _, K, *_ in layers}
    ys = {N: torch.empty((M, N), device=dev, dtype=torch.bfloat16) for _, _, N, *_ in layers}
    step_bytes = sum(bench.algo_bytes(M, K, N) for _, K, N, *_ in layers)

    bufs = [torch.randn((1, 24576), device=dev).to(torch.bfloat16) for _ in range(2)]

    def step():
        for i, (name, K, N, Kl, Nl, mode, lin) in enumerate(layers):
            if args.chain:
                lin.forward_into(bufs[i & 1][:, :K], bufs[(i + 1) & 1][:, :N])
            else:
                lin.forward_into(xs[K], ys[N])
    for cfg in args.cfg or [""]:
        kv = dict(item.split("=", 1) for item in cfg.split(",") if item)
        for k, v in kv.items():
            os.environ[k] = v
        if args.api == "chain":
            if args.chain:
                steps = bench.chain_steps(torch, layers, M, dev)
            else:
                steps = [(lin, xs[K], ys[N]) for name, K, N, Kl, Nl, mode, lin in layers]
            dchain = P.DecodeChain(steps)
            print("[ab] chain kernel:", dchain.kernel(), flush=True)
            step = dchain.run
        best = None
        for rep in range(2):
            st = torch.cuda.Stream(dev)
            with torch.cuda.stream(st):
                step()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=st):
                    step()
                for _ in range(3):
                    g.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(args.steps):
                    g.replay()
                e1.record(st)
                st.synchronize()
            ms = e0.elapsed_time(e1) / args.steps
            best = ms if best is None else min(best, ms)
        print(f"[ab] {cfg or 'default':60s} M={M} {best:.4f} ms/step  {step_bytes / best / 1e6:8.1f} GB/s  "
              f"{best * 1e3 / len(layers):.2f} us/linear", flush=True)
        for k in kv:
            del os.environ[k]


if __name__ == "__main__":
    main()
