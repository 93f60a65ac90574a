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
QWEN3_06B_GATE_UP = ("gate_up", 1024, 6144)                       # BASELINE config 1
QWEN3_4B_GPTQ = [("qkv", 2560, 6144), ("o", 4096, 2560), ("gate_up", 2560, 19456), ("down", 9728, 2560)]   # config 4
SWEEP_M = [16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192]     # config 3
DECODE_M = [1, 2, 4, 8]
GPTQ_M = [1, 2, 4, 8, 16, 32, 64, 128, 256]                       # config 4
GROUP = 128
L2_BYTES = 126e6


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


def decode_traffic(world, chain, n_linears):
    """dram bytes per LINEAR from the committed ncu --set full capture of the kernel in use (profiles/, written together
    with the git hash of the tree it measured), in the unit of `achieved` (per linear, averaged over the step's shapes);
    None for the TP workload (not captured under ncu)."""
    names = ("r2_chain_traffic.json",) if chain else ("r2_decode_traffic.json", "r1_decode_traffic.json")
    for name in names:
        p = os.path.join(ROOT, "profiles", name)
        if world == 1 and os.path.exists(p):
            d = json.load(open(p))
            return d.get("mean_per_launch"), {"file": "profiles/" + name, "git": d.get("git"), "kernel": d.get("kernel")}
    return None, None


def roof_frac(M, K, N, us, peaks):
    """Fraction of the min(HBM, tensor) roofline: the bound is the slower of bytes / HBM peak and flops / bf16 peak."""
    t_hbm = algo_bytes(M, K, N) / (peaks["hbm_gbs"] * 1e9)
    t_tc = 2.0 * M * K * N / (peaks["bf16_tflops"] * 1e12)
    return round(max(t_hbm, t_tc) / (us * 1e-6), 4), ("hbm" if t_hbm >= t_tc else "tensor")


def time_launches(torch, dev, fn, n_inner, target_ms=8.0, max_reps=200):
    """fn() enqueues n_inner launches; captured once in a CUDA graph and replayed: microseconds per launch."""
    st = torch.cuda.Stream(dev)
    with torch.cuda.stream(st):
        fn()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        graph.replay()
        e0.record(st); graph.replay(); e1.record(st)
        st.synchronize()
        one = max(e0.elapsed_time(e1), 1e-3)
        reps = int(max(3, min(max_reps, target_ms / one)))
        e0.record(st)
        for _ in range(reps):
            graph.replay()
        e1.record(st)
        st.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * n_inner)


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
def rand_awq(torch, K, N, dev, gen=None):
    qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev, generator=gen)
    qzeros = torch.randint(-2**31, 2**31 - 1, (K // GROUP, N // 8), dtype=torch.int32, device=dev, generator=gen)
    scales = (torch.rand((K // GROUP, N), device=dev, generator=gen) * 0.018 + 0.002).to(torch.float16)
    return {"qweight": qweight, "qzeros": qzeros, "scales": scales}


def rand_gptq_act_order(torch, K, N, dev):
    """GPTQ checkpoint [K/8, N] with a random act-order g_idx (a permutation of repeat(arange(G), 128))."""
    w = {"qweight": torch.randint(-2**31, 2**31 - 1, (K // 8, N), dtype=torch.int32, device=dev),
         "qzeros": torch.randint(-2**31, 2**31 - 1, (K // GROUP, N // 8), dtype=torch.int32, device=dev),
         "scales": (torch.rand((K // GROUP, N), device=dev) * 0.018 + 0.002).to(torch.float16)}
    g_idx = torch.arange(K, device=dev, dtype=torch.int32) // GROUP
    w["g_idx"] = g_idx[torch.randperm(K, device=dev)].contiguous()
    return w


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
            w = rand_awq(torch, K, N, device, gen)
            if mode == "col":
                lin = P.ColumnParallelW4A16Linear(cfg, K, N, False, comm, device)
            elif mode == "row":
                lin = P.RowParallelW4A16Linear(cfg, K, N, False, comm, device, peers=peers)
            else:
                lin = cfg.create_linear(K, N, False, device)
            lin.load_weights(w)
            layers.append((name, K, N, lin.in_features(), lin.out_features(), mode, lin))
            del w
    torch.cuda.synchronize()
    return layers


def chain_steps(torch, layers, M, dev):
    """(linear, x, y) of every linear of the step with the data flow of the model wherever the shapes allow it: one
    output buffer per shape (re-used layer after layer, as a CUDA-graph replay re-uses its buffers), and every linear
    reads the previous linear's output -- all of it when the widths agree, at M = 1 also a leading part of it
    (qkv 6144 -> o 4096 stands in for attention, gate_up 24576 -> down 12288 for silu*mul); otherwise its own x."""
    ys, xs, steps, prev = {}, {}, [], None
    for name, K, N, Kl, Nl, mode, lin in layers:
        y = ys.setdefault((name, Nl), torch.zeros((M, Nl), device=dev, dtype=torch.bfloat16))
        if prev is not None and prev.shape[1] == Kl:
            x = prev
        elif prev is not None and M == 1 and prev.shape[1] > Kl:
            x = prev[:, :Kl]
        else:
            x = xs.setdefault((name, Kl), (torch.randn((M, Kl), device=dev, dtype=torch.float32) * 0.05).to(torch.bfloat16))
        steps.append((lin, x, y))
        prev = y
    return steps


def timed_steps(torch, dist, dev, run, stream, args, world, rank, local):
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
    return ms / args.steps, clk


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
    tp_chk = tp_check(P, torch, dist, dev, nccl, peers, rank, world) if world > 1 else None
    if tp_chk is not None and not tp_chk["ok"]:
        if rank == 0:
            print(json.dumps({"metric": METRIC, "n_gpus": world, "tp_check": tp_chk, "error": "tp_check failed"}))
        raise SystemExit(3)
    layers = build_layers(P, torch, shapes, n_layers, dev, nccl, peers=peers)

    xs = {}
    ys = {}
    for name, K, N, Kl, Nl, mode, lin in layers:
        xs.setdefault((name, Kl), torch.randn((M, Kl), device=dev, dtype=torch.float32).to(torch.bfloat16))
        ys.setdefault((name, Nl), torch.empty((M, Nl), device=dev, dtype=torch.bfloat16))
    step_bytes = sum(algo_bytes(M, K, N) for _, K, N, *_ in layers)  # whole-job (unsharded) bytes

    def one_step():
        for name, K, N, Kl, Nl, mode, lin in layers:
            lin.forward_into(xs[(name, Kl)], ys[(name, Nl)])  # row-parallel: GEMV + all-reduce inside

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
        ms_per_step, clk = timed_steps(torch, dist, dev, run, stream, args, world, rank, local)
    per_linear = {"api": "w4a16_forward_into per linear, one CUDA graph per step, programmatic launches",
                  "ms_per_step": round(ms_per_step, 4), "GBps": round(step_bytes / (ms_per_step * 1e-3) / 1e9, 1),
                  "kernel": layers[0][6].kernel_for(M), "launches_per_step": int(launches_per_step),
                  "data_flow": "static x per shape (linears ordered by the stream only)"}
    # N = 1 headline: the same 144 linears as ONE DecodeChain (w4a16_chain_run, one persistent launch per step) with the
    # model's data flow -- every linear reads what the previous one wrote, so the order is a true dependency
    chain, chain_kernel, launch = None, None, "cuda-graph replay" if graph is not None else "stream launches"
    if args.api == "chain" and (world == 1 or peers is not None):
        # N > 1: one chain per rank over its shards; the row shards all-reduce inside the launch (w4a16_chain_set_peer).
        # Creation is collective-safe: every rank takes the same branch (same shapes, same peer group state).
        try:
            chain = P.DecodeChain(chain_steps(torch, layers, M, dev), peers=peers if world > 1 else None)
            chain_kernel = chain.kernel()
            if not chain_kernel.startswith("chain_imma"):
                chain = None
        except P.W4A16Error as e:
            if rank == 0:
                print(f"bench: chain unavailable ({e}); per-linear launches", file=sys.stderr)
            chain = None
    if chain is not None:
        cgraph = None
        launches0 = P.kernel_launches()
        with torch.cuda.stream(stream):
            chain.run()
            launches_per_step = P.kernel_launches() - launches0
            if args.graph:
                cgraph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(cgraph, stream=stream):
                    chain.run()
            ms_per_step, clk = timed_steps(torch, dist, dev, cgraph.replay if cgraph is not None else chain.run, stream, args,
                                           world, rank, local)
        per_linear["frac_measured_hbm"] = round(per_linear["GBps"] / peaks["hbm_gbs"], 4)
        launch = "one persistent launch per step (w4a16_chain_run" + (", CUDA-graph replay)" if cgraph is not None else ")")
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
                   "launch": launch, "pdl": os.environ.get("W4A16_NO_PDL", "0") == "0",
                   "data_flow": ("every linear reads the previous linear's output (bench.chain_steps): true dependencies"
                                 if chain is not None else "static x per shape"),
                   "allreduce": (None if world == 1 else
                                 "inside the chain launch: row shards exchange fp32 partials over NVLink peer memory (no NCCL launch)"
                                 if chain is not None else
                                 "fused GEMV epilogue over NVLink peer memory (one kernel per linear, no NCCL launch)"
                                 if peers is not None else "ncclAllReduce after the GEMV")},
        "gpu_launches": int(launches_per_step * args.steps),
        "clocks": clk,
    }
    if tp_chk is not None:
        out["tp_check"] = tp_chk
    # -- roofline of the dominant kernel (decode GEMV) ---------------------------------
    n_launch = 1 if chain is not None else n_lin          # the chain is ONE launch per step
    per_launch_bytes = step_bytes / n_launch / max(world, 1)
    per_launch_ms = ms_per_step / n_launch
    achieved = per_launch_bytes / (per_launch_ms * 1e-3) / 1e9
    traffic, traffic_src = decode_traffic(world, chain is not None, n_lin)
    if chain is not None:
        out["per_linear_api"] = per_linear
    out["roofline"] = {"bound": "hbm", "kernel": chain_kernel if chain is not None else layers[0][6].kernel_for(M),
                       "achieved": round(achieved, 1),
                       "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": round(achieved / peaks["hbm_gbs"], 4),
                       "traffic": traffic, "traffic_source": traffic_src, "peak_source": peaks["source"],
                       "frac_of_spec_8TBs": round(achieved / 8000.0, 4),
                       "algorithmic_bytes_per_launch": int(per_launch_bytes)}
    if rank == 0 and world == 1:
        if chain is not None:
            # the same step with 2 and 4 tokens (multi-token decode): widths that agree are real hand-overs (o -> gate_up,
            # down -> next qkv), the others (qkv -> o, gate_up -> down: the model has attention / silu*mul there) read
            # their own static x and are not ordered against their predecessor -- an upper-side figure, not the headline
            out["chain_m"] = []
            for Mm in (2, 4):
                try:
                    cm = P.DecodeChain(chain_steps(torch, layers, Mm, dev))
                    us = time_launches(torch, dev, cm.run, 1, target_ms=20.0, max_reps=40)
                    sb = sum(algo_bytes(Mm, K, N) for _, K, N, *_ in layers)
                    out["chain_m"].append({"M": Mm, "ms_per_step": round(us / 1e3, 4), "GBps": round(sb / (us * 1e-6) / 1e9, 1),
                                           "frac_measured_hbm": round(sb / (us * 1e-6) / 1e9 / peaks["hbm_gbs"], 4), "kernel": cm.kernel()})
                    del cm
                except P.W4A16Error as e:
                    out["chain_m"].append({"M": Mm, "unsupported": str(e)})
        if not args.quick:
            if chain is not None:
                out["e2e"] = e2e_chain(P, torch, dev, chain, layers, args)
                out["e2e_per_linear"] = e2e_host(P, torch, dev, layers, args)
            else:
                out["e2e"] = e2e_host(P, torch, dev, layers, args)
        del layers, xs, ys, graph, chain
        torch.cuda.empty_cache()
        out["per_shape"], out["sweep"] = shape_sweep(P, torch, dev, peaks, args)
        if not args.quick:
            out["prefill"] = prefill_point(out["sweep"], peaks)
            out["prefill_sustained"] = prefill_sustained(P, torch, dev, peaks, local)
            out["layer_chain"] = layer_chain(P, torch, dev, peaks)
            out["config1"] = config1(P, torch, dev, peaks)
            out["gptq_act_order"] = gptq_act_order(P, torch, dev, peaks)
            out["reference_gpu"] = reference_gpu(torch, dev, peaks)
            out["cpu_baseline"] = cpu_baseline()
    elif chain is not None:
        # end to end at N > 1: every rank feeds its chain from pinned host memory and reads the last linear's output back
        # (w4a16_chain_run_host: H2D, the tensor-parallel chain, D2H, sync); SPMD, wall clock, MAX over ranks
        hx = torch.randn((M, layers[0][3])).mul_(0.05).to(torch.bfloat16).pin_memory()
        hy = torch.empty((M, layers[-1][4]), dtype=torch.bfloat16).pin_memory()
        n_e2e = max(3, min(args.steps, 20))
        chain.run_host(hx, hy)
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            chain.run_host(hx, hy)
        dt = torch.tensor([(time.perf_counter() - t0) / n_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if rank == 0:
            out["e2e"] = {"value": round(step_bytes / float(dt.item()) / 1e9, 1), "unit": "GB/s", "ms_per_step": round(float(dt.item()) * 1e3, 3),
                          "h2d_bytes_per_step": int(hx.numel() * 2) * world, "d2h_bytes_per_step": int(hy.numel() * 2) * world,
                          "api": "w4a16_chain_run_host on every rank (pinned host x in, y of the last linear out, sync per step); max over ranks"}
    elif rank == 0:
        out["e2e"] = {"value": None, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                      "note": "host-buffer path measured at N=1 only"}
    if world > 1 and not args.quick:
        pf = prefill_tp(torch, dist, dev, layers[:len(shapes)], peaks, world)   # every rank takes part (all-reduce inside)
        if rank == 0:
            out["prefill_tp"] = pf
        # the SAME workload on one GPU (rank 0 alone, the others wait): strong-scaling efficiency on one workload
        del layers, xs, ys, graph
        torch.cuda.empty_cache()
        one = same_workload_1gpu(P, torch, dev, shapes, n_layers, args) if rank == 0 else None
        dist.barrier()
        if rank == 0:
            out["same_workload_1gpu_ms"] = one
            out["speedup_vs_1gpu_same_workload"] = round(one / ms_per_step, 3) if one else None
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


def tp_check(P, torch, dist, dev, nccl, peers, rank, world):
    """Before anything is timed: one small row-parallel layer through the fused GEMV + all-reduce epilogue, compared on
    EVERY rank with (a) the CPU oracle, (b) the GEMV + ncclAllReduce path, (c) the other ranks' bits.  The oracle is
    used as the checker only (it never runs inside a timed region)."""
    import numpy as np
    from oracle import oracle as O
    from helpers import awq_unpacked, bits_of, oracle_forward, t_bf16_from_bits, t_i32
    K, N, g = 8192, 8192, GROUP        # Llama-3-70B o_proj
    res = {"shape": f"{K}->{N} row-parallel, TP={world}", "cases": [], "ok": True}
    try:
        qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, g, seed=7)
        q, z, sb = awq_unpacked(qweight, qzeros, scales)
        w = {"qweight": t_i32(qweight).to(dev), "qzeros": t_i32(qzeros).to(dev), "scales": torch.from_numpy(scales).to(dev)}
        cfg = P.AwqConfig.int4(g)
        fused = P.RowParallelW4A16Linear(cfg, K, N, False, nccl, dev, peers=peers)
        fused.load_weights(w)
        plain = P.RowParallelW4A16Linear(cfg, K, N, False, nccl, dev)
        plain.load_weights(w)
        k0, k1 = P.shard_range(K, rank, world)
        for M in (1, 4, 1):
            xb = O.make_activations(M, K, seed=40 + M)
            want = oracle_forward(xb, q, z, sb, g)
            x = t_bf16_from_bits(np.ascontiguousarray(xb[:, k0:k1])).to(dev)
            yf = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=dev)
            yn = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
            fused.forward_into(x, yf)
            plain.forward_into(x, yn)
            torch.cuda.synchronize()
            scale = float(np.abs(want).max())
            ef = float(np.abs(O.np_bf16_bits_to_f32(bits_of(yf)).astype(np.float64) - want).max()) / scale
            en = float(np.abs(O.np_bf16_bits_to_f32(bits_of(yn)).astype(np.float64) - want).max()) / scale
            ys = [torch.empty_like(yf) for _ in range(world)]
            dist.all_gather(ys, yf)
            same = all(torch.equal(o.view(torch.int16), yf.view(torch.int16)) for o in ys)
            # fp32 reduction in the fused path: the 1e-2 bar; the bf16 NCCL reduce (the reference's semantics) grows with world
            ok = np.isfinite(ef) and ef <= 1e-2 and en <= (1e-2 if world <= 2 else 2e-2) and same
            flags = torch.tensor([0 if ok else 1, int(ef * 1e6), int(en * 1e6)], device=dev, dtype=torch.int64)
            dist.all_reduce(flags, op=dist.ReduceOp.MAX)
            res["cases"].append({"M": M, "fused_vs_oracle_max_rel": round(flags[1].item() / 1e6, 6),
                                 "nccl_vs_oracle_max_rel": round(flags[2].item() / 1e6, 6),
                                 "bit_identical_across_ranks": bool(same), "all_ranks_ok": flags[0].item() == 0})
            res["ok"] = res["ok"] and flags[0].item() == 0
        res["fused_path"] = fused.kernel_for(1) if peers is not None else "ncclAllReduce (no peer group)"
        if peers is not None:
            # the chain kernel's in-launch all-reduce: [column shard -> row shard] as one DecodeChain, the row shard's
            # output against the oracle of the full linear applied to the gathered input, and across ranks
            Nc = 1024 * world
            qw2, qz2, sc2, _ = O.make_awq_checkpoint(K, Nc, g, seed=8)
            col = P.ColumnParallelW4A16Linear(cfg, K, Nc, False, nccl, dev)
            col.load_weights({"qweight": t_i32(qw2).to(dev), "qzeros": t_i32(qz2).to(dev), "scales": torch.from_numpy(sc2 * np.float16(0.2)).to(dev)})
            qw3, qz3, sc3, _ = O.make_awq_checkpoint(Nc, N, g, seed=9)
            q3, z3, sb3 = awq_unpacked(qw3, qz3, sc3)
            row = P.RowParallelW4A16Linear(cfg, Nc, N, False, nccl, dev)
            row.load_weights({"qweight": t_i32(qw3).to(dev), "qzeros": t_i32(qz3).to(dev), "scales": torch.from_numpy(sc3).to(dev)})
            xc = t_bf16_from_bits(O.make_activations(1, K, seed=60)).to(dev)
            mid = torch.zeros((1, Nc // world), dtype=torch.bfloat16, device=dev)
            yc = torch.full((1, N), float("nan"), dtype=torch.bfloat16, device=dev)
            ch = P.DecodeChain([(col, xc, mid), (row, mid, yc)], peers=peers)
            for _ in range(4):
                ch.run()
            torch.cuda.synchronize()
            parts = [torch.empty_like(mid) for _ in range(world)]
            dist.all_gather(parts, mid)
            want = oracle_forward(bits_of(torch.cat(parts, dim=1)), q3, z3, sb3, g)
            ec = float(np.abs(O.np_bf16_bits_to_f32(bits_of(yc)).astype(np.float64) - want).max()) / float(np.abs(want).max())
            ys = [torch.empty_like(yc) for _ in range(world)]
            dist.all_gather(ys, yc)
            same = all(torch.equal(o.view(torch.int16), yc.view(torch.int16)) for o in ys)
            flags = torch.tensor([0 if (np.isfinite(ec) and ec <= 1e-2 and same) else 1, int(min(ec, 1e3) * 1e6) if np.isfinite(ec) else 10**9],
                                 device=dev, dtype=torch.int64)
            dist.all_reduce(flags, op=dist.ReduceOp.MAX)
            res["chain"] = {"kernel": ch.kernel(), "row_shard_vs_oracle_max_rel": round(flags[1].item() / 1e6, 6),
                            "bit_identical_across_ranks": bool(same), "all_ranks_ok": flags[0].item() == 0}
            res["ok"] = res["ok"] and flags[0].item() == 0
            del ch
    except Exception as e:   # noqa: BLE001
        res["ok"], res["error"] = False, repr(e)
    return res


def prefill_tp(torch, dist, dev, layer, peaks, world, M=4096, reps=5):
    """BASELINE config 5 at M = 4096: one tensor-parallel Llama-3-70B layer (qkv / gate_up column-parallel, o / down
    row-parallel ending in the all-reduce of M x 8192 bf16), stream launches, CUDA events, MAX over ranks per shape.
    TFLOPS are whole-job (the unsharded flops of the shape over the slowest rank's time)."""
    points, tot_flops, tot_us = [], 0.0, 0.0
    stream = torch.cuda.Stream(dev)
    for name, K, N, Kl, Nl, mode, lin in layer:
        x = torch.randn((M, Kl), device=dev, dtype=torch.float32).mul_(0.05).to(torch.bfloat16)
        y = torch.empty((M, Nl), device=dev, dtype=torch.bfloat16)
        with torch.cuda.stream(stream):
            for _ in range(2):
                lin.forward_into(x, y)
            stream.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(reps):
                lin.forward_into(x, y)
            e1.record(stream)
            stream.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / reps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        us = float(t.item())
        flops = 2.0 * M * K * N
        tot_flops += flops
        tot_us += us
        points.append({"shape": f"{name} {K}->{N}", "parallel": mode, "M": M, "us": round(us, 1),
                       "TFLOPS_aggregate": round(flops / us / 1e6, 1), "TFLOPS_per_gpu": round(flops / us / 1e6 / world, 1),
                       "frac_measured_bf16_per_gpu": round(flops / us / 1e6 / world / peaks["bf16_tflops"], 4)})
        del x, y
    return {"M": M, "tp": world, "layer_TFLOPS_aggregate": round(tot_flops / tot_us / 1e6, 1),
            "layer_TFLOPS_per_gpu": round(tot_flops / tot_us / 1e6 / world, 1), "points": points,
            "allreduce": "ncclAllReduce of M x 8192 bf16 after the row-parallel GEMMs (parallel_layers.rs:295-304)"}


def same_workload_1gpu(P, torch, dev, shapes, n_layers, args):
    """The TP workload UNSHARDED on this one GPU (7.1 GB of weights for the 16-layer Llama-3-70B step): ms per step."""
    layers = build_layers(P, torch, [sh[:3] for sh in shapes], n_layers, dev)
    xs = {K: torch.randn((1, K), device=dev).to(torch.bfloat16) for _, K, *_ in layers}
    ys = {N: torch.empty((1, N), device=dev, dtype=torch.bfloat16) for _, _, N, *_ in layers}

    def step():
        for name, K, N, Kl, Nl, mode, lin in layers:
            lin.forward_into(xs[K], ys[N])
    if args.api == "chain":    # the same API the N > 1 headline uses: one DecodeChain launch per step
        chain = P.DecodeChain(chain_steps(torch, layers, 1, dev))
        if chain.kernel().startswith("chain_imma"):
            step = chain.run
    us = time_launches(torch, dev, step, 1, target_ms=30.0, max_reps=30)
    return round(us / 1e3, 4)


def shape_sweep(P, torch, dev, peaks, args):
    """Each Qwen3-8B shape alone: decode M in {1,2,4,8} and BASELINE config 3's M = 16 ... 8192.  Small M rotates over
    enough weight copies to exceed 2x L2 (the weights are the traffic); from M = 512 on the activations alone exceed L2.
    Every point carries its fraction of the min(HBM, tensor) roofline."""
    cfg = P.AwqConfig.int4(GROUP)
    decode, sweep = [], []
    ms_list = DECODE_M + (SWEEP_M[:3] if args.quick else SWEEP_M)
    for name, K, N in QWEN3_8B:
        wbytes = algo_bytes(1, K, N)
        copies = max(2, int(2 * L2_BYTES // wbytes) + 1)
        lins = []
        for c in range(copies):
            lin = cfg.create_linear(K, N, False, dev)
            lin.load_weights(rand_awq(torch, K, N, dev))
            lins.append(lin)
        for M in ms_list:
            use = lins if M < 512 else lins[:2]
            x = torch.randn((M, K), device=dev).to(torch.bfloat16)
            y = torch.empty((M, N), device=dev, dtype=torch.bfloat16)

            def fn():
                for lin in use:
                    lin.forward_into(x, y)
            us = time_launches(torch, dev, fn, len(use))
            gbs = algo_bytes(M, K, N) / (us * 1e-6) / 1e9
            tf = 2.0 * M * K * N / (us * 1e-6) / 1e12
            frac, bound = roof_frac(M, K, N, us, peaks)
            row = {"shape": f"{name} {K}->{N}", "M": M, "us": round(us, 2), "GBps": round(gbs, 1), "TFLOPS": round(tf, 1),
                   "roofline_frac": frac, "bound": bound, "kernel": use[0].kernel_for(M), "copies": len(use)}
            if M in DECODE_M:
                row["frac_measured_hbm"] = round(gbs / peaks["hbm_gbs"], 3)
                decode.append(row)
            else:
                sweep.append(row)
            del x, y
        del lins
        torch.cuda.empty_cache()
    return decode, sweep


def prefill_point(sweep, peaks):
    """BASELINE's prefill headline: the four Qwen3-8B shapes at M = 4096, aggregate TFLOPS (from the sweep's points)."""
    pts = [r for r in sweep if r["M"] == 4096]
    if not pts:
        return None
    flops = sum(r["TFLOPS"] * r["us"] for r in pts)      # TFLOPS * us = MFLOP
    agg = flops / sum(r["us"] for r in pts)
    return {"M4096_TFLOPS": round(agg, 1), "M4096_frac_measured_bf16": round(agg / peaks["bf16_tflops"], 4),
            "M4096_frac_spec_2250": round(agg / 2250.0, 4), "peak": peaks["bf16_tflops"],
            "timing": "burst: a few ms of back-to-back launches per shape (see prefill_sustained for the seconds-long figure)",
            "points": pts}


def prefill_sustained(P, torch, dev, peaks, gpu_index, seconds=2.5):
    """M = 4096 on the four Qwen3-8B shapes looped for >= `seconds` with the clock sampler running: the figure to set
    against bf16_tflops_sustained (cuBLAS looped for 4 s under the 1 kW power cap)."""
    cfg = P.AwqConfig.int4(GROUP)
    M = 4096
    work = []
    for name, K, N in QWEN3_8B:
        lin = cfg.create_linear(K, N, False, dev)
        lin.load_weights(rand_awq(torch, K, N, dev))
        work.append((lin, torch.randn((M, K), device=dev).to(torch.bfloat16), torch.empty((M, N), device=dev, dtype=torch.bfloat16)))
    flops_pass = sum(2.0 * M * K * N for _, K, N in QWEN3_8B)

    def one_pass():
        for lin, x, y in work:
            lin.forward_into(x, y)
    for _ in range(3):
        one_pass()
    torch.cuda.synchronize()
    clocks = ClockSampler(gpu_index)
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0, passes = time.perf_counter(), 0
    e0.record()
    while time.perf_counter() - t0 < seconds:
        for _ in range(20):
            one_pass()
        passes += 20
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop()
    tf = flops_pass * passes / (ms * 1e-3) / 1e12
    return {"M": M, "seconds": round(ms / 1e3, 2), "passes": passes, "TFLOPS": round(tf, 1),
            "frac_measured_bf16_sustained": round(tf / peaks["bf16_tflops_sustained"], 4),
            "frac_spec_2250": round(tf / 2250.0, 4), "peak_sustained": peaks["bf16_tflops_sustained"], "clocks": clk}


def config1(P, torch, dev, peaks):
    """BASELINE config 1: one Qwen3-0.6B gate_up linear, 1024 -> 6144, M = 1: the GPU kernel over rotating weight
    copies (> 2x L2) and the reference's CPU path on this box, single-thread (faithful: awq.rs:142-161 is a scalar
    loop) and all cores, median of 3."""
    name, K, N = QWEN3_06B_GATE_UP
    cfg = P.AwqConfig.int4(GROUP)
    copies = int(2 * L2_BYTES // algo_bytes(1, K, N)) + 1
    lins = []
    for c in range(copies):
        lin = cfg.create_linear(K, N, False, dev)
        lin.load_weights(rand_awq(torch, K, N, dev))
        lins.append(lin)
    x = torch.randn((1, K), device=dev).to(torch.bfloat16)
    y = torch.empty((1, N), device=dev, dtype=torch.bfloat16)

    def fn():
        for lin in lins:
            lin.forward_into(x, y)
    us = time_launches(torch, dev, fn, copies)
    gbs = algo_bytes(1, K, N) / (us * 1e-6) / 1e9
    out = {"shape": f"{name} {K}->{N}", "M": 1, "gpu_us": round(us, 2), "gpu_GBps": round(gbs, 1),
           "gpu_frac_measured_hbm": round(gbs / peaks["hbm_gbs"], 3), "kernel": lins[0].kernel_for(1), "copies": copies}
    out["cpu"] = cpu_time_shapes([(K, N)], reps=3)
    return out


def layer_chain(P, torch, dev, peaks, n_layers=36, Ms=(1, 2, 4)):
    """SURVEY 8f-1 + 8f-4 together: the Qwen3-8B decoder layer MINUS attention as the engine would chain it.  Per layer
    [o_proj (x = the attention output, a static buffer), gate/up as ONE SwiGLU handle with the post-attention residual
    add + RMSNorm as its prologue, down_proj, next layer's fused qkv with residual add + input RMSNorm as its prologue];
    attention itself sits between qkv and o (it is not part of this library), so every layer is one DecodeChain and a
    step is n_layers launches (one CUDA graph).  Same weight bytes as the headline step; the residual stream, the norms
    and silu*mul never leave the kernel."""
    H, I, QKV = 4096, 12288, 6144
    cfg = P.AwqConfig.int4(GROUP)
    gen = torch.Generator(device=dev)
    gen.manual_seed(99)
    layers = []
    for li in range(n_layers):
        o = cfg.create_linear(H, H, False, dev)
        o.load_weights(rand_awq(torch, H, H, dev, gen))
        gu = cfg.create_linear(H, 2 * I, False, dev)
        gu.set_epilogue(P.EPILOGUE_SWIGLU)
        for half in range(2):
            gu.load_part(half * I, rand_awq(torch, H, I, dev, gen))
        down = cfg.create_linear(I, H, False, dev)
        down.load_weights(rand_awq(torch, I, H, dev, gen))
        qkv = cfg.create_linear(H, QKV, False, dev)
        qkv.load_weights(rand_awq(torch, H, QKV, dev, gen))
        w_post = (torch.rand(H, device=dev, generator=gen) + 0.5).to(torch.bfloat16)
        w_in = (torch.rand(H, device=dev, generator=gen) + 0.5).to(torch.bfloat16)
        layers.append((o, gu, down, qkv, w_post, w_in))
    step_bytes = n_layers * sum(algo_bytes(1, K, N) for _, K, N in QWEN3_8B)
    out = {"workload": f"{n_layers} x [o, gate_up+SwiGLU (post-attention add + RMSNorm prologue), down, qkv (add + input RMSNorm "
                       "prologue)], one DecodeChain launch per layer, attention output and residual stream as static inputs",
           "points": []}
    for M in Ms:
        z = lambda n: torch.zeros((M, n), device=dev, dtype=torch.bfloat16)
        attn = (torch.randn((M, H), device=dev) * 0.05).to(torch.bfloat16)
        resid = (torch.randn((M, H), device=dev) * 0.05).to(torch.bfloat16)
        y_o, y_gu, y_dn, y_qkv, r_mid, r_out = z(H), z(I), z(H), z(QKV), z(H), z(H)
        chains = []
        try:
          for o, gu, down, qkv, w_post, w_in in layers:
            chains.append(P.DecodeChain(
                [(o, attn, y_o), (gu, y_o, y_gu), (down, y_gu, y_dn), (qkv, y_dn, y_qkv)],
                prologues=[None, {"norm_weight": w_post, "residual_in": resid, "residual_out": r_mid, "eps": 1e-6},
                           None, {"norm_weight": w_in, "add_running_residual": True, "residual_out": r_out, "eps": 1e-6}]))
        except P.W4A16Error as e:      # e.g. M = 4: four tokens x K = 12288 staged + the residual leave no room for the ring
            out["points"].append({"M": M, "unsupported": str(e)})
            del chains
            continue
        kern = chains[0].kernel()

        def fn():
            for c in chains:
                c.run()
        us = time_launches(torch, dev, fn, 1, target_ms=20.0, max_reps=40)
        gbs = step_bytes / (us * 1e-6) / 1e9
        out["points"].append({"M": M, "ms_per_step": round(us / 1e3, 4), "GBps": round(gbs, 1), "kernel": kern,
                              "frac_measured_hbm": round(gbs / peaks["hbm_gbs"], 4), "launches_per_step": n_layers})
        del chains
    return out


def gptq_act_order(P, torch, dev, peaks):
    """BASELINE config 4: Qwen3-4B GPTQ int4 with an act-order g_idx, M = 1 ... 256 (the activation gather of the
    act-order path is inside the timed launches)."""
    cfg = P.GptqConfig(4, GROUP, desc_act=True)
    rows = []
    for name, K, N in QWEN3_4B_GPTQ:
        copies = max(2, int(2 * L2_BYTES // algo_bytes(1, K, N)) + 1)
        lins = []
        for c in range(copies):
            lin = cfg.create_linear(K, N, False, dev)
            lin.load_weights(rand_gptq_act_order(torch, K, N, dev))
            lin.reserve(max(GPTQ_M))
            lins.append(lin)
        for M in GPTQ_M:
            x = torch.randn((M, K), device=dev).to(torch.bfloat16)
            y = torch.empty((M, N), device=dev, dtype=torch.bfloat16)

            def fn():
                for lin in lins:
                    lin.forward_into(x, y)
            us = time_launches(torch, dev, fn, copies)
            frac, bound = roof_frac(M, K, N, us, peaks)
            rows.append({"shape": f"{name} {K}->{N}", "M": M, "us": round(us, 2),
                         "GBps": round(algo_bytes(M, K, N) / (us * 1e-6) / 1e9, 1),
                         "TFLOPS": round(2.0 * M * K * N / (us * 1e-6) / 1e12, 1), "roofline_frac": frac, "bound": bound,
                         "kernel": lins[0].kernel_for(M)})
        del lins
        torch.cuda.empty_cache()
    # the whole decode step of the model as one DecodeChain: 36 layers x (qkv, o, gate_up, down), every linear gathering
    # its act-order x from the previous linear's mailbox inside the launch; the same step as per-linear launches beside it
    step = {"layers": 36, "points": []}
    layers = []
    for li in range(36):
        for name, K, N in QWEN3_4B_GPTQ:
            lin = cfg.create_linear(K, N, False, dev)
            lin.load_weights(rand_gptq_act_order(torch, K, N, dev))
            lin.reserve(4)
            layers.append((name, K, N, K, N, None, lin))
    for M in (1, 2, 4):
        sb = sum(algo_bytes(M, K, N) for _, K, N, *_ in layers)
        steps = chain_steps(torch, layers, M, dev)
        pt = {"M": M}
        try:
            ch = P.DecodeChain(steps)
            us = time_launches(torch, dev, ch.run, 1, target_ms=20.0, max_reps=40)
            pt.update({"chain_ms_per_step": round(us / 1e3, 4), "chain_GBps": round(sb / (us * 1e-6) / 1e9, 1),
                       "chain_frac_measured_hbm": round(sb / (us * 1e-6) / 1e9 / peaks["hbm_gbs"], 4), "kernel": ch.kernel()})
            del ch
        except P.W4A16Error as e:
            pt["chain_unsupported"] = str(e)

        def per_linear():
            for lin, x, y in steps:
                lin.forward_into(x, y)
        us = time_launches(torch, dev, per_linear, 1, target_ms=20.0, max_reps=40)
        pt.update({"per_linear_ms_per_step": round(us / 1e3, 4), "per_linear_GBps": round(sb / (us * 1e-6) / 1e9, 1)})
        step["points"].append(pt)
    del layers
    torch.cuda.empty_cache()
    return {"workload": "Qwen3-4B GPTQ int4 g128, act-order g_idx (random permutation of groups)", "points": rows, "decode_step": step}


def reference_gpu(torch, dev, peaks):
    """The kernels to beat: the REFERENCE'S OWN CUDA kernels, compiled where they lie into oracle/_ref
    (checker-side library, never loaded by the product), timed on this box outside every product region:
    awq_gemv_int4_kt_bf16 at M = 1 (awq_gemv.cu:571-665, launch geometry of marlin_cuda.rs:572-602) and, for prefill,
    awq_marlin_dequant_int4_bf16 + a dense bf16 matmul at M = 4096 (awq_marlin_dequant.cu:35-69 followed by cuBLAS
    through torch.matmul, as marlin_cuda.rs:793-818 does through candle)."""
    import ctypes
    import numpy as np
    so = os.path.join(ROOT, "oracle", "_ref", "libref_kernels.so")
    if not os.path.exists(so):
        return {"unavailable": "oracle/_ref/libref_kernels.so not built (needs /root/reference at build time)"}
    from oracle import oracle as O
    from helpers import t_bf16_from_bits, t_i32
    ref = ctypes.CDLL(so)
    vp, i = ctypes.c_void_p, ctypes.c_int
    ref.ref_awq_marlin_dequant_int4_bf16.argtypes = [vp, vp, vp, vp, i, i, i, vp]
    ref.ref_awq_gemv_int4_kt_bf16.argtypes = [vp, vp, vp, vp, vp, vp, i, i, i, i, vp]
    rows = []
    for name, K, N in QWEN3_8B:
        qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, GROUP, seed=K ^ N)
        # the reference's resident layout (awq_marlin.rs:621-656), built by the oracle's restatement of its repacks
        qw_t = t_i32(O.C.transpose_u32(O.C.awq_to_gptq_qweight(qweight, K, N))).to(dev)
        qz = t_i32(O.C.repack_awq_nibbles(qzeros)).to(dev)
        sb = t_bf16_from_bits(O.np_f32_to_bf16_bits(np.asarray(scales, np.float32))).to(dev)
        copies = max(2, int(2 * L2_BYTES // algo_bytes(1, K, N)) + 1)
        sets = [(qw_t.clone(), qz.clone(), sb.clone()) for _ in range(copies)]   # distinct addresses: no L2 reuse
        x = torch.randn((1, K), device=dev).to(torch.bfloat16)
        y = torch.empty((1, N), device=dev, dtype=torch.bfloat16)

        def gemv():
            s = torch.cuda.current_stream().cuda_stream
            for a, b, c in sets:
                ref.ref_awq_gemv_int4_kt_bf16(y.data_ptr(), x.data_ptr(), a.data_ptr(), c.data_ptr(), b.data_ptr(), None, 1, N, K, GROUP, s)
        us = time_launches(torch, dev, gemv, copies)
        gbs = algo_bytes(1, K, N) / (us * 1e-6) / 1e9
        M = 4096
        xp = torch.randn((M, K), device=dev).to(torch.bfloat16)
        W = torch.empty((K, N), dtype=torch.bfloat16, device=dev)
        yp = torch.empty((M, N), dtype=torch.bfloat16, device=dev)

        def prefill():
            s = torch.cuda.current_stream().cuda_stream
            ref.ref_awq_marlin_dequant_int4_bf16(W.data_ptr(), qw_t.data_ptr(), sb.data_ptr(), qz.data_ptr(), K, N, GROUP, s)
            torch.matmul(xp, W, out=yp)
        usp = time_launches(torch, dev, prefill, 1)
        rows.append({"shape": f"{name} {K}->{N}", "gemv_M1_us": round(us, 2), "gemv_M1_GBps": round(gbs, 1),
                     "gemv_M1_frac_measured_hbm": round(gbs / peaks["hbm_gbs"], 3),
                     "dequant_matmul_M4096_us": round(usp, 1),
                     "dequant_matmul_M4096_TFLOPS": round(2.0 * M * K * N / (usp * 1e-6) / 1e12, 1)})
        del sets, W, xp, yp
        torch.cuda.empty_cache()
    tot_b = sum(algo_bytes(1, K, N) for _, K, N in QWEN3_8B)
    tot_f = sum(2.0 * 4096 * K * N for _, K, N in QWEN3_8B)
    return {"what": "the reference's own kernels (oracle/_ref, compiled from /root/reference/crates/core/kernels) on this GPU",
            "gemv_M1_layer_GBps": round(tot_b / sum(r["gemv_M1_us"] for r in rows) / 1e3, 1),
            "dequant_matmul_M4096_layer_TFLOPS": round(tot_f / sum(r["dequant_matmul_M4096_us"] for r in rows) / 1e6, 1),
            "points": rows}


def e2e_chain(P, torch, dev, chain, layers, args):
    """The end-to-end decode step through the C ABI with HOST buffers: w4a16_chain_run_host = H2D copy of the step's
    input (the first linear's x, pinned), the chain, D2H copy of the step's result (the last linear's y), stream sync.
    The activations between the linears stay on the device, as they do in the reference."""
    M = 1
    hx = torch.randn((M, layers[0][3])).mul_(0.05).to(torch.bfloat16).pin_memory()
    hy = torch.empty((M, layers[-1][4]), dtype=torch.bfloat16).pin_memory()
    steps = max(3, min(args.steps, 20))
    chain.run_host(hx, hy)
    t0 = time.perf_counter()
    for _ in range(steps):
        chain.run_host(hx, hy)
    dt = (time.perf_counter() - t0) / steps
    step_bytes = sum(algo_bytes(M, K, N) for _, K, N, *_ in layers)
    return {"value": round(step_bytes / dt / 1e9, 1), "unit": "GB/s", "ms_per_step": round(dt * 1e3, 3),
            "h2d_bytes_per_step": int(hx.numel() * 2), "d2h_bytes_per_step": int(hy.numel() * 2),
            "api": "w4a16_chain_run_host per step (pinned host x of the first linear in, y of the last linear out, sync per step)"}


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
def cpu_time_shapes(shapes, reps=3, modes=("single_thread", "all_cores")):
    """Reference CPU path (AwqLinear::forward, awq.rs:522-567: scalar AWQ dequant on EVERY call + fp32 GEMM) over the
    given (K, N) list at M = 1: median of `reps`, single thread (faithful) and all host cores."""
    import numpy as np
    from oracle import oracle as O
    O.build()
    cases = []
    for K, N in shapes:
        qweight, qzeros, scales, _ = O.make_awq_checkpoint(K, N, GROUP, seed=K ^ N)
        x = O.make_activations(1, K, seed=1)
        scratch = (np.empty((K, N), np.float32), np.empty((1, K), np.float32))
        cases.append((K, N, qweight, qzeros, scales.astype(np.float32), x, scratch))
    nbytes = sum(algo_bytes(1, K, N) for K, N in shapes)
    out = {}
    for label, threads in (("single_thread", 1), ("all_cores", O.C.max_threads())):
        if label not in modes:
            continue
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            for K, N, qweight, qzeros, sf, x, scratch in cases:
                O.C.awq_linear_forward(x, qweight, sf, qzeros, None, GROUP, K, N, threads=threads, scratch=scratch)
            ts.append(time.perf_counter() - t0)
        ts.sort()
        med = ts[len(ts) // 2]
        out[label] = {"GBps": round(nbytes / med / 1e9, 4), "ms": round(med * 1e3, 2), "threads": threads, "reps": reps,
                      "min_ms": round(ts[0] * 1e3, 2), "max_ms": round(ts[-1] * 1e3, 2)}
    return out


def cpu_baseline():
    """cpu_baseline of the N=1 line: one Qwen3-8B layer (4 linears, M=1) of the headline workload through the oracle's
    restatement of AwqLinear::forward on this box's host cores; `value` is the all-cores median (the fairer figure),
    the faithful single-thread figure sits beside it."""
    r = cpu_time_shapes([(K, N) for _, K, N in QWEN3_8B], reps=3)
    ac, st = r["all_cores"], r["single_thread"]
    return {"value": ac["GBps"], "unit": "GB/s", "cores": ac["threads"], "kind": "port",
            "single_thread": st, "all_cores": ac,
            "sample": f"1 of 36 Qwen3-8B layers (4 linears, M=1): scalar AWQ dequant on every call + fp32 GEMM "
                      f"(awq.rs:522-567 restated in oracle/w4a16_oracle.c); median of 3: {ac['ms']} ms on {ac['threads']} "
                      f"threads, {st['ms']} ms on 1 thread (the reference's own loop is single-threaded)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # each step = one Qwen3-8B layer sample on all host cores (bounded: the whole run ends within minutes)
    import numpy as np  # noqa: F401
    from oracle import oracle as O
    O.build()
    th = O.C.max_threads()
    shapes = [(K, N) for _, K, N in QWEN3_8B]
    nbytes = sum(algo_bytes(1, K, N) for K, N in shapes)
    for _ in range(min(args.warmup, 1)):
        cpu_time_shapes(shapes, reps=1, modes=("all_cores",))
    vals = []
    for _ in range(max(1, min(args.steps, 3))):
        r = cpu_time_shapes(shapes, reps=1, modes=("all_cores",))["all_cores"]
        vals.append(r["ms"] / 1e3)
    secs = sorted(vals)[len(vals) // 2]
    gbs = nbytes / secs / 1e9
    sample = (f"1 of 36 Qwen3-8B layers per step (4 linears, M=1), {th} host threads, median of {len(vals)} steps; "
              "reference = AwqLinear::forward restated (Rust toolchain absent: the crate cannot be built here)")
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
    ap.add_argument("--api", default="chain", choices=["chain", "linear"],
                    help="N=1 headline: the step as one DecodeChain (persistent launch) or as per-linear launches")
    ap.add_argument("--no-fused-ar", dest="fused_ar", action="store_false",
                    help="TP runs: all-reduce through ncclAllReduce instead of the fused GEMV epilogue over peer memory")
    ap.add_argument("--quick", action="store_true", help="decode step + short sweep only (no prefill / e2e / cpu legs)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    run_gpu(args)


if __name__ == "__main__":
    main()
