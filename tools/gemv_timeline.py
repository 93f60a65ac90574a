#!/usr/bin/env python
"""Per-phase timeline of the decode kernel inside a PDL chain (W4A16_GV_DEBUG stamps).
    W4A16_GV_DEBUG=1 python tools/gemv_timeline.py [--shape qkv] [--m 1] [--graph]
Prints, for the last 3 launches of a chain of `--n` launches over distinct weight copies, the
min / median / max over CTAs of every stamp, relative to the earliest start stamp of the first of them."""
import argparse
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
SHAPES = {"qkv": (4096, 6144), "o": (4096, 4096), "gate_up": (4096, 24576), "down": (12288, 4096)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="qkv")
    ap.add_argument("--m", type=int, default=1)
    ap.add_argument("--n", type=int, default=16)
    ap.add_argument("--graph", action="store_true")
    args = ap.parse_args()
    os.environ["W4A16_GV_DEBUG"] = "1"
    import numpy as np
    import torch
    import conftest
    P = conftest._load_product()
    dev = torch.device("cuda", 0)
    K, N = SHAPES[args.shape]
    cfg = P.AwqConfig.int4(128)
    lins = []
    for c in range(args.n):
        qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev)
        qzeros = torch.randint(-2**31, 2**31 - 1, (K // 128, N // 8), dtype=torch.int32, device=dev)
        scales = (torch.rand((K // 128, N), device=dev) * 0.018 + 0.002).to(torch.float16)
        lin = cfg.create_linear(K, N, False, dev)
        lin.load_weights({"qweight": qweight, "qzeros": qzeros, "scales": scales})
        lins.append(lin)
    x = torch.randn((args.m, K), device=dev).to(torch.bfloat16)
    y = torch.empty((args.m, N), device=dev, dtype=torch.bfloat16)
    st = torch.cuda.Stream(dev)
    with torch.cuda.stream(st):
        for lin in lins:
            lin.forward_into(x, y)
        st.synchronize()
        if args.graph:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                for lin in lins:
                    lin.forward_into(x, y)
            g.replay()
            g.replay()
        else:
            for lin in lins:
                lin.forward_into(x, y)
        st.synchronize()
    words = 4 * 1024 * 8
    buf = (ctypes.c_uint64 * words)()
    n = ctypes.c_uint64()
    rc = P.lib().w4a16_debug_stamps(ctypes.cast(buf, ctypes.c_void_p), words, ctypes.byref(n))
    assert rc == 0, P.lib().w4a16_last_error()
    a = np.frombuffer(buf, dtype=np.uint64).reshape(4, 1024, 8).astype(np.int64)
    last = int(n.value) - 1
    order = [(last - 2) % 4, (last - 1) % 4, last % 4]
    ncta = int((a[order[0], :256, 0] > 0).sum())
    t0 = a[order[0], :ncta, 0].min()
    names = ["start", "dep-wait", "staged", "streamed", "combined", "end", "stage0-in", "last-in"]
    print(f"shape {args.shape} K={K} N={N} M={args.m} kernel={lins[0].kernel_for(args.m)} ctas={ncta} launches={n.value}")
    for j, slot in enumerate(order):
        print(f"-- launch {j - 2:+d}")
        for i, nm in enumerate(names):
            if a[slot, :ncta, i].max() == 0:
                continue
            v = (a[slot, :ncta, i] - t0) / 1e3
            print(f"   {nm:9s} min {v.min():8.2f}  med {np.median(v):8.2f}  max {v.max():8.2f} us")
        if j == 2:
            trace(a, slot)


def trace(a, slot):
    tr = a[slot].reshape(-1)[2048:2048 + 16 * 8].reshape(16, 8)
    if tr.max() == 0:
        return
    t0 = tr[0, 0]
    print("   warp-0 checkpoints of CTA 7 (cycles): top, waited, loads-in, refill-issued, imma-done, end")
    for j in range(16):
        if tr[j, 0] == 0:
            break
        print("    unit", j, " ".join(f"{int(v - t0):7d}" for v in tr[j, :6]))


if __name__ == "__main__":
    main()
