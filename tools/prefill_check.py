#!/usr/bin/env python
"""Quick prefill check/timing on the GPU box: product forward vs torch fp32 matmul on the product's own
(bit-exact-checked elsewhere) dequantised weights.  Not a test: a development probe.
    python tools/prefill_check.py [--time] [--shapes qkv o ...] [--m 64 256 4096]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

SHAPES = {"k4096": (4096, 256), "k4096n1024": (4096, 1024), "k1024": (1024, 256), "k2048": (2048, 512), "n6144": (512, 6144), "tiny": (256, 128), "tail": (512, 192), "qkv": (4096, 6144), "o": (4096, 4096),
          "gate_up": (4096, 24576), "down": (12288, 4096)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, nargs="*", default=[33, 64, 100, 128, 256, 300, 1024, 4096])
    ap.add_argument("--shapes", nargs="*", default=["tiny", "tail", "qkv"])
    ap.add_argument("--g", type=int, default=128)
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--bias", action="store_true")
    ap.add_argument("--dispatch", default="auto")
    args = ap.parse_args()
    import torch
    import conftest
    P = conftest._load_product()
    P.set_dispatch(args.dispatch)
    dev = torch.device("cuda", 0)
    cfg = P.AwqConfig.int4(args.g)
    g = args.g
    bad = 0
    for name in args.shapes:
        K, N = SHAPES[name]
        G = K // g if g > 0 else 1
        qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev)
        qzeros = torch.randint(-2**31, 2**31 - 1, (G, N // 8), dtype=torch.int32, device=dev)
        scales = (torch.rand((G, N), device=dev) * 0.018 + 0.002).to(torch.float16)
        lin = cfg.create_linear(K, N, args.bias, dev)
        w = {"qweight": qweight, "qzeros": qzeros, "scales": scales}
        if args.bias:
            w["bias"] = torch.randn((N,), device=dev).to(torch.bfloat16)
        lin.load_weights(w)
        W = lin.dequantize().float()
        for M in args.m:
            x = torch.randn((M, K), device=dev).to(torch.bfloat16)
            y = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16)
            lin.forward_into(x, y)
            torch.cuda.synchronize()
            ref = x.float() @ W
            if args.bias:
                ref = ref + w["bias"].float()
            err = (y.float() - ref).abs().max().item()
            scale = ref.abs().max().item()
            ok = err <= 1e-2 * scale and not torch.isnan(y.float()).any().item()
            bad += 0 if ok else 1
            line = f"{name} K={K} N={N} M={M} kernel={lin.kernel_for(M)} max_err={err:.4g} max|y|={scale:.4g} {'ok' if ok else 'FAIL'}"
            if args.time:
                for _ in range(3):
                    lin.forward_into(x, y)
                torch.cuda.synchronize()
                reps = 10
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(reps):
                    lin.forward_into(x, y)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / reps
                line += f" {ms*1e3:.1f} us {2.0*M*K*N/ms/1e9:.1f} TFLOPS"
            print(line, flush=True)
    print("FAILED" if bad else "ALL OK", bad)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
