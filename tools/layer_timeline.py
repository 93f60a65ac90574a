#!/usr/bin/env python
"""Per-phase timeline of the decode kernels inside the REAL step mix (qkv, o, gate_up, down chained by programmatic
launches in one CUDA graph), from the W4A16_GV_DEBUG per-CTA %globaltimer stamps.
    python tools/layer_timeline.py [--m 1] [--layers 6] [--chain]
Prints the last layer's four launches: min / median / max over CTAs of every stamp, relative to the moment the first of
them (qkv) saw its dependency resolved.  --chain: every linear reads what the previous one wrote (M = 1 only)."""
import argparse
import ctypes
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
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--chain", action="store_true")
    args = ap.parse_args()
    os.environ["W4A16_GV_DEBUG"] = "1"
    import numpy as np
    import torch
    import conftest
    P = conftest._load_product()
    dev = torch.device("cuda", 0)
    cfg = P.AwqConfig.int4(128)
    lins = []
    for c in range(args.layers):
        for name, K, N in SHAPES:
            qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev)
            qzeros = torch.randint(-2**31, 2**31 - 1, (K // 128, N // 8), dtype=torch.int32, device=dev)
            scales = (torch.rand((K // 128, N), device=dev) * 0.018 + 0.002).to(torch.float16)
            lin = cfg.create_linear(K, N, False, dev)
            lin.load_weights({"qweight": qweight, "qzeros": qzeros, "scales": scales})
            lins.append((name, lin))
    xs = {K: torch.randn((args.m, K), device=dev).to(torch.bfloat16) for _, K, _ in SHAPES}
    ys = {N: torch.empty((args.m, N), device=dev, dtype=torch.bfloat16) for _, _, N in SHAPES}
    if args.chain:
        assert args.m == 1
        big = torch.randn((1, 24576), device=dev).to(torch.bfloat16)
        bufs = [big.clone() for _ in range(2)]

    def step():
        for i, (name, lin) in enumerate(lins):
            if args.chain:
                lin.forward_into(bufs[i & 1][:, :lin.in_features()], bufs[(i + 1) & 1][:, :lin.out_features()])
            else:
                lin.forward_into(xs[lin.in_features()], ys[lin.out_features()])

    st = torch.cuda.Stream(dev)
    with torch.cuda.stream(st):
        step()
        st.synchronize()
        if not args.no_graph:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                step()
            for _ in range(3):
                g.replay()
        else:
            step()
        st.synchronize()
    words = 4 * 1024 * 8
    buf = (ctypes.c_uint64 * words)()
    n = ctypes.c_uint64()
    rc = P.lib().w4a16_debug_stamps(ctypes.cast(buf, ctypes.c_void_p), words, ctypes.byref(n))
    assert rc == 0, P.lib().w4a16_last_error()
    a = np.frombuffer(buf, dtype=np.uint64).reshape(4, 1024, 8).astype(np.int64)
    last = int(n.value) - 1
    order = [(last - 3 + j) % 4 for j in range(4)]
    names = ["start", "dep-wait", "staged", "streamed", "combined", "end", "x-in/st0", "last-in"]
    t0 = None
    prev_end = None
    for j, slot in enumerate(order):
        ncta = int((a[slot, :256, 0] > 0).sum())
        if t0 is None:
            t0 = a[slot, :ncta, 1].min()
        nm_l, lin = lins[len(lins) - 4 + j]
        print(f"-- {nm_l} kernel={lin.kernel_for(args.m)} ctas={ncta}")
        for i, nm in enumerate(names):
            if a[slot, :ncta, i].max() == 0:
                continue
            v = (a[slot, :ncta, i] - t0) / 1e3
            print(f"   {nm:9s} min {v.min():8.2f}  med {np.median(v):8.2f}  max {v.max():8.2f} us")
        end = (a[slot, :ncta, 5].max() - t0) / 1e3
        if prev_end is not None:
            print(f"   => kernel-to-kernel {end - prev_end:6.2f} us")
        prev_end = end


if __name__ == "__main__":
    main()
