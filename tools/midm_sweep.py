#!/usr/bin/env python
"""Mid-M sweep on the GPU box: each shape alone, CUDA-graph replay over rotating weight copies (> 2x L2),
for a list of M and a forced dispatch ("auto", "decode", "prefill").  A development probe, not a test.
    python tools/midm_sweep.py --m 8 16 32 64 128 256 --dispatch auto prefill [--shapes qkv o gate_up down]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

SHAPES = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 24576), "down": (12288, 4096),
          "l70_tp2_qkv": (8192, 5120), "l70_tp2_o": (4096, 8192), "l70_tp2_gate_up": (8192, 28672), "l70_tp2_down": (14336, 8192),
          "l70_tp8_qkv": (8192, 1280), "l70_tp8_o": (1024, 8192), "l70_tp8_gate_up": (8192, 7168), "l70_tp8_down": (3584, 8192),
          "q4_qkv": (2560, 6144), "q4_o": (4096, 2560), "q4_gate_up": (2560, 19456), "q4_down": (9728, 2560)}
G = 128


def algo_bytes(M, K, N, g=G):
    return K * N // 2 + (K // g) * N * 2 + (K // g) * N // 2 + 2 * M * K + 2 * M * N


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, nargs="*", default=[8, 16, 32, 64, 128, 256])
    ap.add_argument("--shapes", nargs="*", default=["qkv", "o", "gate_up", "down"])
    ap.add_argument("--dispatch", nargs="*", default=["auto"])
    ap.add_argument("--check", action="store_true", help="also compare with torch fp32 on dequantised weights")
    args = ap.parse_args()
    import torch
    import conftest
    P = conftest._load_product()
    dev = torch.device("cuda", 0)
    cfg = P.AwqConfig.int4(G)
    for name in args.shapes:
        K, N = SHAPES[name]
        copies = max(2, int(2 * 126e6 // algo_bytes(1, K, N)) + 1)
        lins = []
        for c in range(copies):
            qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev)
            qzeros = torch.randint(-2**31, 2**31 - 1, (K // G, N // 8), dtype=torch.int32, device=dev)
            scales = (torch.rand((K // G, N), device=dev) * 0.018 + 0.002).to(torch.float16)
            lin = cfg.create_linear(K, N, False, dev)
            lin.load_weights({"qweight": qweight, "qzeros": qzeros, "scales": scales})
            lins.append(lin)
        W = lins[0].dequantize().float() if args.check else None
        torch.cuda.synchronize()   # the side stream below must not reuse the freed bf16 temporary while .float() still reads it
        for M in args.m:
            x = torch.randn((M, K), device=dev).to(torch.bfloat16)
            y = torch.empty((M, N), device=dev, dtype=torch.bfloat16)
            for disp in args.dispatch:
                P.set_dispatch(disp)
                st = torch.cuda.Stream(dev)
                with torch.cuda.stream(st):
                    for lin in lins:
                        lin.forward_into(x, y)
                    st.synchronize()
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph, stream=st):
                        for lin in lins:
                            lin.forward_into(x, y)
                    reps = max(3, 300 // copies)
                    graph.replay()
                    st.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(st)
                    for _ in range(reps):
                        graph.replay()
                    e1.record(st)
                    st.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / (reps * copies)
                gbs = algo_bytes(M, K, N) / (us * 1e-6) / 1e9
                tf = 2.0 * M * K * N / (us * 1e-6) / 1e12
                line = f"{name} K={K} N={N} M={M} dispatch={disp} kernel={lins[0].kernel_for(M)} {us:.2f} us {gbs:.0f} GB/s {tf:.1f} TF"
                if args.check:
                    yy = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16)
                    lins[0].forward_into(x, yy)
                    torch.cuda.synchronize()
                    ref = x.float() @ W
                    err = (yy.float() - ref).abs().max().item()
                    sc = ref.abs().max().item()
                    line += f" max_err={err:.3g}/{sc:.3g} {'ok' if err <= 1e-2 * sc else 'FAIL'}"
                print(line, flush=True)
            P.set_dispatch("auto")
        del lins


if __name__ == "__main__":
    main()
