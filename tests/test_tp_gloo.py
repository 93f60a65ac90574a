"""World-size-2 CPU test (gloo) of the tensor-parallel host logic (parallel_layers.rs:36-324):
shard ranges, checkpoint slicing for column / row shards, the all-reduce(Sum) + bias order of
RowParallelLinear, through vllm_rust_b200.parallel's TorchDistCommunicator.  The per-shard GEMM is the
ORACLE here (no GPU in this container); the same composition runs on CUDA shards in
tests/test_parity_gpu.py::test_tp_shards_compose and across real ranks in bench.py --gpus N."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_q):
    try:
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        import conftest
        P = conftest._load_product()
        from oracle import oracle as O
        from helpers import awq_unpacked, oracle_forward
        comm = P.TorchDistCommunicator()
        assert comm.rank() == rank and comm.world_size() == world
        K, N, g, M = 1024, 512, 128, 3
        qweight, qzeros, scales, b = O.make_awq_checkpoint(K, N, g, seed=5, bias=True)   # same on every rank
        q, z, sb = awq_unpacked(qweight, qzeros, scales)
        xb = O.make_activations(M, K, seed=9)
        bias = O.np_bf16_round(b)
        want = oracle_forward(xb, q, z, sb, g, bias_f32=bias)
        # column parallel: rank r owns columns [n0, n1); AWQ words hold 8 consecutive outputs, so the slice is word aligned
        n0, n1 = P.shard_range(N, rank, world)
        assert n0 % 8 == 0 and (n1 - n0) % 64 == 0
        qs, zs, ss = awq_unpacked(qweight[:, n0 // 8:n1 // 8], qzeros[:, n0 // 8:n1 // 8], scales[:, n0:n1])
        y_col = oracle_forward(xb, qs, zs, ss, g, bias_f32=bias[n0:n1])
        gathered = [torch.zeros((M, n1 - n0), dtype=torch.float64) for _ in range(world)]
        dist.all_gather(gathered, torch.from_numpy(np.ascontiguousarray(y_col)))
        assert np.allclose(torch.cat(gathered, dim=1).numpy(), want, rtol=0, atol=1e-9)
        # row parallel: rank r owns rows [k0, k1) and group rows [k0/g, k1/g); partial -> all_reduce(Sum) -> + bias
        k0, k1 = P.shard_range(K, rank, world)
        assert (k1 - k0) % g == 0
        part = oracle_forward(xb[:, k0:k1], q[k0:k1], z[k0 // g:k1 // g], sb[k0 // g:k1 // g], g)
        t = torch.from_numpy(np.ascontiguousarray(part))
        comm.all_reduce(t)
        y_row = t.numpy() + bias[None, :]
        assert np.abs(y_row - want).max() <= 1e-6 * max(1.0, np.abs(want).max())
        # the shape gate applies to the shard (SURVEY 8d gate check)
        P.check_marlin_supports_shape(n1 - n0, K, g)
        P.check_marlin_supports_shape(N, k1 - k0, g)
        try:
            P.shard_range(N + 1, rank, world)
            raise AssertionError("shard_range accepted an indivisible dimension")
        except ValueError:
            pass
        # the fused all-reduce's peer buffers need a GPU: without one the collective constructor fails on EVERY rank
        # alike (no hang, no CPU fallback), so a caller can fall back to NCCL consistently
        try:
            P.PeerGroup(rank, world, 0, 1024)
            raise AssertionError("PeerGroup succeeded without a GPU")
        except P.W4A16Error as e:
            assert "PeerGroup" in str(e)
        dist.barrier()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        import traceback
        out_q.put((rank, traceback.format_exc()))
        return
    out_q.put((rank, None))


def test_tp_world_size_2_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, err in results:
        assert err is None, f"rank {rank}:\n{err}"
