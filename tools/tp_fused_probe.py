#!/usr/bin/env python
"""Under torchrun (N ranks): one row-parallel Llama-70B o_proj / down_proj shard, M tokens; device time per call of
(a) the local GEMV alone, (b) GEMV + ncclAllReduce, (c) the fused GEMV + all-reduce epilogue; stream launches and
CUDA-graph replay.  A development probe.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/tp_fused_probe.py [--m 1]"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=1)
    ap.add_argument("--reps", type=int, default=200)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import conftest
    P = conftest._load_product()
    comm = P.NcclCommunicator(rank, world, local)
    peers = P.PeerGroup(rank, world, local, 8192)
    cfg = P.AwqConfig.int4(128)
    for name, K, N in [("o", 8192, 8192), ("down", 28672, 8192)]:
        gen = torch.Generator(device=dev); gen.manual_seed(7)
        qweight = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device=dev, generator=gen)
        qzeros = torch.randint(-2**31, 2**31 - 1, (K // 128, N // 8), dtype=torch.int32, device=dev, generator=gen)
        scales = (torch.rand((K // 128, N), device=dev, generator=gen) * 0.018 + 0.002).to(torch.float16)
        w = {"qweight": qweight, "qzeros": qzeros, "scales": scales}
        copies = 6   # rotate weights beyond L2
        fused = [P.RowParallelW4A16Linear(cfg, K, N, False, comm, dev, peers=peers) for _ in range(copies)]
        plain = [P.RowParallelW4A16Linear(cfg, K, N, False, comm, dev) for _ in range(copies)]
        for l in fused + plain:
            l.load_weights(w)
        Kl = fused[0].in_features()
        x = torch.randn((args.m, Kl), device=dev).to(torch.bfloat16)
        y = torch.empty((args.m, N), device=dev, dtype=torch.bfloat16)
        variants = {"local": lambda i: plain[i % copies].inner.forward_into(x, y),
                    "nccl": lambda i: plain[i % copies].forward_into(x, y),
                    "fused": lambda i: fused[i % copies].forward_into(x, y)}
        st = torch.cuda.Stream(dev)
        for vname, fn in variants.items():
            for mode in ("stream", "graph"):
                if mode == "graph" and vname == "nccl":
                    continue
                with torch.cuda.stream(st):
                    for i in range(copies):
                        fn(i)
                    st.synchronize()
                    dist.barrier()
                    if mode == "graph":
                        g = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(g, stream=st):
                            for i in range(copies * 4):
                                fn(i)
                        run, per = g.replay, copies * 4
                    else:
                        run, per = (lambda: [fn(i) for i in range(copies * 4)]), copies * 4
                    run(); st.synchronize(); dist.barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    n = max(1, args.reps // per)
                    e0.record(st)
                    for _ in range(n):
                        run()
                    e1.record(st)
                    st.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / (n * per)
                t = torch.tensor([us], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
                if rank == 0:
                    print(f"{name} K_local={Kl} N={N} M={args.m} tp={world} {vname:6s} {mode:6s} {t.item():8.2f} us/call", flush=True)
        del fused, plain
    dist.barrier()
    peers.destroy(); comm.destroy(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
