#!/usr/bin/env python
"""bench.py -- W4A16 linear on B200: decode GB/s (M=1) and prefill TFLOPS (M=4096).

Workloads (BASELINE.json configs):
  N=1   configs[1]: Qwen3-8B AWQ int4 g128, all 36 decoder layers x the four fused
        projection shapes (qkv 4096->6144, o 4096->4096, gate_up 4096->24576,
        down 12288->4096), M=1.  One step = one decode pass over the 144 linears
        = 3.6 GB of distinct weights (28x the 126 MB L2, so no step re-reads L2).
        The same shapes at M=4096 give the prefill TFLOPS (configs[2] point).
  N>1   configs[4]: Llama-3-70B-shaped AWQ int4 layers, tensor-parallel over N ranks
        (qkv/gate_up column-parallel, o/down row-parallel + NCCL all-reduce), M=1.
        Fixed total work -> "strong".

`value` is device-resident throughput (CUDA events around K steps); `e2e` goes
through w4a16_forward_host with pinned HOST buffers (H2D + kernel + D2H + sync per
linear).  `--impl reference` times the reference's CPU path (the oracle's restatement
of AwqLinear::forward, awq.rs:522-567: scalar dequant on every call + fp32 GEMM).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "W4A16 linear: decode GB/s (M=1) and prefill TFLOPS (M=4096), % of B200 roofline"
QWEN3_8B = [("qkv", 4096, 6144), ("o", 4096, 4096), ("gate_up", 4096, 24576), ("down", 12288, 4096)]
QWEN3_8B_LAYERS = 36
LLAMA3_70B = [("qkv", 8192, 10240, "col"), ("o", 8192, 8192, "row"), ("gate_up", 8192, 57344, "col"),
              ("down", 28672, 8192, "row")]
GROUP = 128


def algo_bytes(M, K, N, g=GROUP):
    """int4 weights + 16-bit scales + int4 zeros + bf16 x + bf16 y."""
    return K * N // 2 + (K // g) * N * 2 + (K // g) * N // 2 + 2 * M * K + 2 * M * N


def load_product():
    import conftest  # tests/conftest.py owns the hyphenated-package import
    return conftest._load_product()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def decode_traffic(world):
    """dram bytes per launch of the decode kernel from the committed ncu --set full capture (profiles/), averaged
    over the four Qwen3-8B shapes like `achieved`; None for the TP workload (not captured under ncu)."""
    p = os.path.join(ROOT, "profiles", "r1_decode_traffic.json")
    if world != 1 or not os.path.exists(p):
        return None
    return json.load(open(p)).get("mean_per_launch")


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.proc, self.lines, self.idx = None, [], gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx = float(f[2])
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# -------------------------------------------------------------------------------------
def build_layers(P, torch, shapes, n_layers, device, comm=None, seed=0, peers=None):
    """Synthetic random-int4 AWQ checkpoints of the given shapes, generated on the device and
    loaded through the product's loader API (repack included).  With a communicator every rank
    generates the same FULL checkpoint tensor and its Column/RowParallel layer keeps its slice."""
    gen = torch.Generator(device=device)
    gen.manual_seed(1234 + seed)
    cfg = P.AwqConfig.int4(GROUP)
    layers = []
    for li in range(n_layers):
        for sh in shapes:
            name, K, N = sh[0], sh[1], sh[2]
            mode = sh[3] if len(sh) > 3 and comm is not None else None
            qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=device, generator=gen)
            qzeros = torch.randint(-2**31, 2**31 - 1, (K // GROUP, N // 8), dtype=torch.int32, device=device, generator=gen)
            scales = (torch.rand((K // GROUP, N), device=device, generator=gen) * 0.018 + 0.002).to(torch.float16)
            if mode == "col":
                lin = P.ColumnParallelW4A16Linear(cfg, K, N, False, comm, device)
            elif mode == "row":
                lin = P.RowParallelW4A16Linear(cfg, K, N, False, comm, device, peers=peers)
            else:
                lin = cfg.create_linear(K, N, False, device)
            lin.load_weights({"qweight": qweight, "qzeros": qzeros, "scales": scales})
            layers.append((name, K, N, lin.in_features(), lin.out_features(), mode, lin))
            del qweight, qzeros, scales
    torch.cuda.synchronize()
    return layers


def run_gpu(args):
    import torch
    import torch.distributed as dist
    P = load_product()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = measured_peaks()
    M = 1

    if world == 1:
        shapes, n_layers, workload = QWEN3_8B, args.layers or QWEN3_8B_LAYERS, \
            f"Qwen3-8B AWQ int4 g128, {args.layers or QWEN3_8B_LAYERS} layers x (qkv,o,gate_up,down), M=1 decode"
        parallelism = "single"
    else:
        shapes, n_layers = LLAMA3_70B, args.layers or 16
        workload = (f"Llama-3-70B-shaped AWQ int4 g128, {n_layers} layers x (qkv,o,gate_up,down), M=1 decode, "
                    f"TP={world} (col: qkv/gate_up; row+all-reduce: o/down)")
        parallelism = f"tp{world}"
    nccl = P.NcclCommunicator(rank, world, local) if world > 1 else None
    peers = None
    if world > 1 and args.fused_ar:   # symmetric NVLink buffers: the row-parallel GEMVs all-reduce in their own epilogue
        try:
            peers = P.PeerGroup(rank, world, local, max(sh[2] for sh in shapes if len(sh) > 3 and sh[3] == "row"))
        except P.W4A16Error as e:   # raised on every rank alike: fall back to ncclAllReduce, and say so in `config`
            if rank == 0:
                print(f"bench: peer buffers unavailable ({e}); using ncclAllReduce", file=sys.stderr)
            peers = None
    layers = build_layers(P, torch, shapes, n_layers, dev, nccl, peers=peers)

    xs = {}
    ys = {}
    for name, K, N, Kl, Nl, mode, lin in layers:
        xs.setdefault((name, Kl), torch.randn((M, Kl), device=dev, dtype=torch.float32).to(torch.bfloat16))
        ys.setdefault((name, Nl), torch.empty((M, Nl), device=dev, dtype=torch.bfloat16))
    step_bytes = sum(algo_bytes(M, K, N) for _, K, N, *_ in layers)  # whole-job (unsharded) bytes

    def one_step():
        for name, K, N, Kl, Nl, mode, lin in layers:
            lin.forward_into(xs[(name, Kl)], ys[(name, Nl)])  # row-parallel: GEMV + ncclAllReduce inside

    stream = torch.cuda.Stream(dev)
    launches0 = P.kernel_launches()
    graph = None
    with torch.cuda.stream(stream):
        one_step()  # warm-up outside capture (also first-touch of every weight)
        launches_per_step = P.kernel_launches() - launches0
        # TP with NCCL all-reduces: stream launches (replaying a graph with captured NCCL hung on this box); with the fused
        # epilogue there is no NCCL call in the step and the whole TP step is one CUDA graph
        if args.graph and (world == 1 or (peers is not None and os.environ.get("W4A16_NO_FUSED_AR", "0") == "0")):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream):
                one_step()
        run = graph.replay if graph is not None else one_step
        for _ in range(args.warmup):
            run()
        stream.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        clocks = ClockSampler(local)
        if rank == 0:
            clocks.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            run()
        e1.record(stream)
        stream.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        clk = clocks.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    gbs = step_bytes / (ms_per_step * 1e-3) / 1e9
    n_lin = len(layers)

    out = {
        "metric": METRIC, "value": round(gbs, 1), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
        "scaling": "weak" if world == 1 else "strong", "vs_baseline": None, "dtype": "int4 weights x bf16, fp32 accumulate",
        "data": "synthetic (uniform random int4 nibbles and zeros, f16 scales U(0.002,0.02), bf16 N(0,1) activations)",
        "config": {"workload": workload, "group_size": GROUP, "M": M, "parallelism": parallelism,
                   "linears_per_step": n_lin, "weight_bytes_per_step": step_bytes,
                   "l2_policy": "inputs larger than L2: every step streams %.2f GB of distinct weights" % (step_bytes / 1e9 / max(world, 1)),
                   "launch": "cuda-graph replay" if graph is not None else "stream launches", "pdl": os.environ.get("W4A16_NO_PDL", "0") == "0",
                   "allreduce": (None if world == 1 else "fused GEMV epilogue over NVLink peer memory (one kernel, no NCCL launch)"
                                 if peers is not None else "ncclAllReduce after the GEMV")},
        "gpu_launches": int(launches_per_step * args.steps),
        "clocks": clk,
    }
    # -- roofline of the dominant kernel (decode GEMV) ---------------------------------
    per_launch_bytes = step_bytes / n_lin / max(world, 1)
    per_launch_ms = ms_per_step / n_lin
    achieved = per_launch_bytes / (per_launch_ms * 1e-3) / 1e9
    out["roofline"] = {"bound": "hbm", "kernel": layers[0][6].kernel_for(M), "achieved": round(achieved, 1),
                       "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": round(achieved / peaks["hbm_gbs"], 4),
                       "traffic": decode_traffic(world), "peak_source": peaks["source"],
                       "frac_of_spec_8TBs": round(achieved / 8000.0, 4),
                       "algorithmic_bytes_per_launch": int(per_launch_bytes)}
    if rank == 0 and world == 1:
        out["per_shape"] = per_shape_decode(P, torch, dev, peaks, args)
        if not args.quick:
            out["prefill"] = prefill_point(P, torch, dev, peaks, args)
            out["e2e"] = e2e_host(P, torch, dev, layers, args)
            out["cpu_baseline"] = cpu_baseline(sample_layers=1, threads=None, reps=1)
    elif rank == 0:
        out["e2e"] = {"value": None, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                      "note": "host-buffer path measured at N=1 only"}
    if world > 1 and not args.quick:
        pf = prefill_tp(torch, dist, dev, layers[:len(shapes)], peaks, world)   # every rank takes part (all-reduce inside)
        if rank == 0:
            out["prefill_tp"] = pf
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        if peers is not None:
            peers.destroy()
        if nccl:
            nccl.destroy()
        dist.destroy_process_group()


def per_shape_decode(P, torch, dev, peaks, args):
    """Each Qwen3-8B shape alone, M in {1,8,16}: rotate over enough weight copies to exceed 2x L2."""
    res = []
    cfg = P.AwqConfig.int4(GROUP)
    for name, K, N in QWEN3_8B:
        wbytes = algo_bytes(1, K, N)
        copies = max(2, int(2 * 126e6 // wbytes) + 1)
        lins = []
        for c in range(copies):
            qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev)
            qzeros = torch.randint(-2**31, 2**31 - 1, (K // GROUP, N // 8), dtype=torch.int32, device=dev)
            scales = (torch.rand((K // GROUP, N), device=dev) * 0.018 + 0.002).to(torch.float16)
            lin = cfg.create_linear(K, N, False, dev)
            lin.load_weights({"qweight": qweight, "qzeros": qzeros, "scales": scales})
            lins.append(lin)
        for M in (1, 8, 16, 32):
            x = torch.randn((M, K), device=dev).to(torch.bfloat16)
            y = torch.empty((M, N), device=dev, dtype=torch.bfloat16)
            st = torch.cuda.Stream(dev)
            with torch.cuda.stream(st):
                for lin in lins:
                    lin.forward_into(x, y)
                graph = torch.cuda.CUDAGraph()  # one pass over all copies; replayed: no host launch cost in the timing
                with torch.cuda.graph(graph, stream=st):
                    for lin in lins:
                        lin.forward_into(x, y)
                reps = max(3, 400 // copies)
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
            res.append({"shape": f"{name} {K}->{N}", "M": M, "us": round(us, 2), "GBps": round(gbs, 1),
                        "frac_measured_hbm": round(gbs / peaks["hbm_gbs"], 3), "kernel": lins[0].kernel_for(M),
                        "copies": copies})
        del lins
    return res


def prefill_point(P, torch, dev, peaks, args):
    """Qwen3-8B shapes at M=4096 (and a short sweep): TFLOPS against the dense bf16 peak."""
    cfg = P.AwqConfig.int4(GROUP)
    rows = []
    lins = {}
    for name, K, N in QWEN3_8B:
        qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev)
        qzeros = torch.randint(-2**31, 2**31 - 1, (K // GROUP, N // 8), dtype=torch.int32, device=dev)
        scales = (torch.rand((K // GROUP, N), device=dev) * 0.018 + 0.002).to(torch.float16)
        lin = cfg.create_linear(K, N, False, dev)
        lin.load_weights({"qweight": qweight, "qzeros": qzeros, "scales": scales})
        lins[name] = lin
    total_flops, total_ms = 0.0, 0.0
    for M in args.prefill_m:
        for name, K, N in QWEN3_8B:
            lin = lins[name]
            x = torch.randn((M, K), device=dev).to(torch.bfloat16)
            y = torch.empty((M, N), device=dev, dtype=torch.bfloat16)
            for _ in range(2):
                lin.forward_into(x, y)
            torch.cuda.synchronize()
            reps = 5
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                lin.forward_into(x, y)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            tf = 2.0 * M * K * N / (ms * 1e-3) / 1e12
            rows.append({"shape": f"{name} {K}->{N}", "M": M, "ms": round(ms, 4), "TFLOPS": round(tf, 1),
                         "frac_measured_bf16": round(tf / peaks["bf16_tflops"], 3), "kernel": lin.kernel_for(M)})
            if M == 4096:
                total_flops += 2.0 * M * K * N
                total_ms += ms
    agg = total_flops / (total_ms * 1e-3) / 1e12 if total_ms else None
    return {"M4096_TFLOPS": round(agg, 1) if agg else None,
            "M4096_frac_measured_bf16": round(agg / peaks["bf16_tflops"], 4) if agg else None,
            "M4096_frac_spec_2250": round(agg / 2250.0, 4) if agg else None, "peak": peaks["bf16_tflops"],
            "points": rows}


def prefill_tp(torch, dist, dev, layer, peaks, world, M=4096, reps=5):
    """BASELINE config 5 at M=4096: one Llama-3-70B-shaped layer (qkv, o, gate_up, down), tensor-parallel; the
    row-parallel o / down finish with ncclAllReduce of the [M, 8192] bf16 output (64 MB).  Device time, max over ranks."""
    bufs = []
    for name, K, N, Kl, Nl, mode, lin in layer:
        bufs.append((torch.randn((M, Kl), device=dev).to(torch.bfloat16), torch.empty((M, Nl), device=dev, dtype=torch.bfloat16)))
    rows = []
    tot_ms, tot_flops = 0.0, 0.0
    for (name, K, N, Kl, Nl, mode, lin), (x, y) in zip(layer, bufs):
        for _ in range(2):
            lin.forward_into(x, y)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            lin.forward_into(x, y)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        flops = 2.0 * M * K * N          # whole-job (unsharded) flops
        tot_ms += ms
        tot_flops += flops
        rows.append({"shape": f"{name} {K}->{N}", "mode": mode, "ms": round(ms, 4), "TFLOPS": round(flops / (ms * 1e-3) / 1e12, 1)})
    agg = tot_flops / (tot_ms * 1e-3) / 1e12
    return {"M": M, "TFLOPS": round(agg, 1), "TFLOPS_per_gpu": round(agg / world, 1),
            "frac_measured_bf16_per_gpu": round(agg / world / peaks["bf16_tflops"], 4), "points": rows,
            "allreduce": "ncclAllReduce (bf16, M x 8192) after the row-parallel GEMMs"}


def e2e_host(P, torch, dev, layers, args):
    """The reference-facing call with HOST buffers: w4a16_forward_host per linear
    (pinned x -> H2D -> kernel -> D2H -> pinned y -> stream sync), every linear of the step."""
    M = 1
    hx, hy = {}, {}
    for name, K, N, Kl, Nl, mode, lin in layers:
        hx.setdefault(Kl, torch.randn((M, Kl)).to(torch.bfloat16).pin_memory())
        hy.setdefault(Nl, torch.empty((M, Nl), dtype=torch.bfloat16).pin_memory())
        lin.reserve(M)
    steps = max(1, min(args.steps, 5))
    for name, K, N, Kl, Nl, mode, lin in layers:  # warm-up
        lin.forward_host(hx[Kl], hy[Nl])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        for name, K, N, Kl, Nl, mode, lin in layers:
            lin.forward_host(hx[Kl], hy[Nl])
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    step_bytes = sum(algo_bytes(M, K, N) for _, K, N, *_ in layers)
    return {"value": round(step_bytes / dt / 1e9, 1), "unit": "GB/s", "ms_per_step": round(dt * 1e3, 3),
            "h2d_bytes_per_step": int(sum(2 * M * Kl for _, _, _, Kl, *_ in layers)),
            "d2h_bytes_per_step": int(sum(2 * M * Nl for _, _, _, _, Nl, *_ in layers)),
            "api": "w4a16_forward_host per linear (pinned host x/y, sync per call)"}


# -------------------------------------------------------------------------------------
def cpu_run_sample(sample_layers, threads, reps):
    """Reference CPU path (AwqLinear::forward, awq.rs:522-567) on `sample_layers` Qwen3-8B layers."""
    import numpy as np
    from oracle import oracle as O
    O.build()
    threads = threads or O.C.max_threads()
    cases = []
    for name, K, N in QWEN3_8B:
        qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, GROUP, seed=K ^ N)
        x = O.make_activations(1, K, seed=1)
        scratch = (np.empty((K, N), np.float32), np.empty((1, K), np.float32))
        cases.append((K, N, qweight, qzeros, scales.astype(np.float32), x, scratch))
    nbytes = sum(algo_bytes(1, K, N) for K, N, *_ in cases) * sample_layers
    best = None
    for _ in range(reps):
        t0 = time.perf_counter()
        for _ in range(sample_layers):
            for K, N, qweight, qzeros, sf, x, scratch in cases:
                O.C.awq_linear_forward(x, qweight, sf, qzeros, None, GROUP, K, N, threads=threads, scratch=scratch)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return nbytes / best / 1e9, best, threads


def cpu_baseline(sample_layers=1, threads=None, reps=1):
    gbs, secs, th = cpu_run_sample(sample_layers, threads, reps)
    return {"value": round(gbs, 3), "unit": "GB/s", "cores": th, "kind": "port",
            "sample": f"{sample_layers} of 36 Qwen3-8B layers (4 linears, M=1): scalar AWQ dequant on every call + fp32 GEMM "
                      f"(awq.rs:522-567 restated in oracle/w4a16_oracle.c), {secs:.2f} s"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # warm-up W and K timed steps; each step = one layer sample (bounded: whole run within minutes)
    for _ in range(min(args.warmup, 1)):
        cpu_run_sample(1, None, 1)
    vals = []
    for _ in range(min(args.steps, 3)):
        gbs, secs, th = cpu_run_sample(1, None, 1)
        vals.append((gbs, secs))
    gbs = sum(v[0] for v in vals) / len(vals)
    secs = sum(v[1] for v in vals) / len(vals)
    sample = (f"1 of 36 Qwen3-8B layers per step (4 linears, M=1), {th} host threads; reference = AwqLinear::forward "
              "restated (Rust toolchain absent: the crate cannot be built here)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(gbs, 3), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": len(vals), "warmup": min(args.warmup, 1), "ms_per_step": round(secs * 1e3, 2), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int4 weights x bf16, fp32 accumulate", "data": "synthetic",
        "config": {"workload": "Qwen3-8B AWQ int4 g128, (qkv,o,gate_up,down), M=1 decode, CPU", "group_size": GROUP, "M": 1},
        "cpu_baseline": {"value": round(gbs, 3), "unit": "GB/s", "cores": th, "kind": "port", "sample": sample},
        "e2e": {"value": round(gbs, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--layers", type=int, default=0, help="override the number of decoder layers (debug)")
    ap.add_argument("--no-graph", dest="graph", action="store_false")
    ap.add_argument("--no-fused-ar", dest="fused_ar", action="store_false",
                    help="TP runs: all-reduce through ncclAllReduce instead of the fused GEMV epilogue over peer memory")
    ap.add_argument("--prefill-m", type=int, nargs="*", default=[64, 256, 1024, 4096])
    ap.add_argument("--quick", action="store_true", help="skip prefill / e2e / cpu legs")
    ap.add_argument("--only-shapes", action="store_true", help="per-shape decode sweep only")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    run_gpu(args)


if __name__ == "__main__":
    main()
