"""Two real ranks on two GPUs: the fused GEMV + all-reduce epilogue over NVLink peer memory
(w4a16_forward_row_fused) against the oracle, against the NCCL path, and across ranks.
Skipped on a box with fewer than two GPUs (the single-GPU parity tests cover the kernels themselves)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_q):
    try:
        import torch.distributed as dist
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        import conftest
        P = conftest._load_product()
        from oracle import oracle as O
        from helpers import awq_unpacked, oracle_forward, assert_close_ref, t_bf16_from_bits, bits_of, t_i32
        dev = torch.device("cuda", rank)
        K, N, g = 4096, 4096, 128   # K / 8 = 512 per rank at world 8: still whole groups
        qweight, qzeros, scales, b = O.make_awq_checkpoint(K, N, g, seed=31, bias=True)   # same on every rank
        q, z, sb = awq_unpacked(qweight, qzeros, scales)
        bias = O.np_bf16_round(b)
        comm = P.NcclCommunicator(rank, world, rank)
        peers = P.PeerGroup(rank, world, rank, N)
        cfg = P.AwqConfig.int4(g)
        w = {"qweight": t_i32(qweight).to(dev), "qzeros": t_i32(qzeros).to(dev),
             "scales": torch.from_numpy(scales).to(dev), "bias": torch.from_numpy(b).to(torch.bfloat16).to(dev)}
        fused = P.RowParallelW4A16Linear(cfg, K, N, True, comm, dev, peers=peers)
        fused.load_weights(w)
        plain = P.RowParallelW4A16Linear(cfg, K, N, True, comm, dev)
        plain.load_weights(w)
        k0, k1 = P.shard_range(K, rank, world)
        for it, M in enumerate([1, 2, 3, 4, 1, 4, 2, 1, 1, 3, 4, 4]):   # many launches: epoch parity, flag reuse
            xb = O.make_activations(M, K, seed=50 + it)
            want = oracle_forward(xb, q, z, sb, g, bias_f32=bias)
            x = t_bf16_from_bits(np.ascontiguousarray(xb[:, k0:k1])).to(dev)
            y = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=dev)
            fused.forward_into(x, y)
            torch.cuda.synchronize()
            assert_close_ref(bits_of(y), want, f"fused rank {rank} M={M} it={it}")
            y2 = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
            plain.forward_into(x, y2)
            torch.cuda.synchronize()
            # the NCCL path rounds every rank's partial sum to bf16 and reduces in bf16 (the reference's semantics,
            # nccl.rs:1338-1364): its error grows with the world size, so beyond 2 ranks it is held to 2e-2, not 1e-2
            if world <= 2:
                assert_close_ref(bits_of(y2), want, f"nccl rank {rank} M={M}")
            else:
                err2 = np.abs(O.np_bf16_bits_to_f32(bits_of(y2)).astype(np.float64) - want).max()
                assert err2 <= 2e-2 * np.abs(want).max(), f"nccl rank {rank} M={M}: {err2}"
            # every rank holds bit-identical results (sources are summed in rank order everywhere)
            ys = [torch.empty_like(y) for _ in range(world)]
            dist.all_gather(ys, y)
            for other in ys:
                assert torch.equal(other.view(torch.int16), y.view(torch.int16))
        # the launch counter is device-resident: a captured graph of fused calls replays correctly, many times
        M = 2
        xb = O.make_activations(M, K, seed=77)
        want = oracle_forward(xb, q, z, sb, g, bias_f32=bias)
        x = t_bf16_from_bits(np.ascontiguousarray(xb[:, k0:k1])).to(dev)
        y = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
        st = torch.cuda.Stream(dev)
        with torch.cuda.stream(st):
            fused.forward_into(x, y)
            st.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=st):
                for _ in range(3):
                    fused.forward_into(x, y)
            for _ in range(5):
                y.fill_(float("nan"))
                graph.replay()
                st.synchronize()
                assert_close_ref(bits_of(y), want, f"graph replay rank {rank}")
        dist.barrier()
        # ONE PeerGroup shared by row-parallel linears of different N and M, interleaved big, small, small (period 3 =
        # the buffer rotation: the big linear always lands on the same buffer and only small launches run between
        # its uses -- stale words of rows [N_small, N_big) would be summed silently if a launch left any behind)
        Nb = 8192
        peers2 = P.PeerGroup(rank, world, rank, Nb)
        qwb, qzb, scb, _ = O.make_awq_checkpoint(K, Nb, g, seed=32)
        qb_, zb_, sbb = awq_unpacked(qwb, qzb, scb)
        big = P.RowParallelW4A16Linear(cfg, K, Nb, False, comm, dev, peers=peers2)
        big.load_weights({"qweight": t_i32(qwb).to(dev), "qzeros": t_i32(qzb).to(dev), "scales": torch.from_numpy(scb).to(dev)})
        small = P.RowParallelW4A16Linear(cfg, K, N, True, comm, dev, peers=peers2)
        small.load_weights(w)
        for it in range(9):
            for lin, (qq, zz, ss, bb, nn), M in ((big, (qb_, zb_, sbb, None, Nb), 4 if it % 2 else 1),
                                                 (small, (q, z, sb, bias, N), 2), (small, (q, z, sb, bias, N), 1)):
                xb = O.make_activations(M, K, seed=300 + 3 * it + M)
                x = t_bf16_from_bits(np.ascontiguousarray(xb[:, k0:k1])).to(dev)
                y = torch.full((M, nn), float("nan"), dtype=torch.bfloat16, device=dev)
                lin.forward_into(x, y)
                torch.cuda.synchronize()
                assert_close_ref(bits_of(y), oracle_forward(xb, qq, zz, ss, g, bias_f32=bb), f"shared group it={it} N={nn} M={M}")
        dist.barrier()
        peers2.destroy()
        # M > 4 falls back to the NCCL path inside the same entry point
        xb = O.make_activations(8, K, seed=99)
        x = t_bf16_from_bits(np.ascontiguousarray(xb[:, k0:k1])).to(dev)
        y = torch.empty((8, N), dtype=torch.bfloat16, device=dev)
        fused.forward_into(x, y)
        torch.cuda.synchronize()
        assert_close_ref(bits_of(y), oracle_forward(xb, q, z, sb, g, bias_f32=bias), "fallback M=8")
        dist.barrier()
        peers.destroy()
        comm.destroy()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        import traceback
        out_q.put((rank, traceback.format_exc()))
        return
    out_q.put((rank, None))


def _n_gpus():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_row_parallel_allreduce_real_ranks(world):
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    try:
        results = [q.get(timeout=240) for _ in range(world)]
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():   # a rank that failed leaves its peer waiting: do not let it linger
                p.terminate()
    for rank, err in results:
        assert err is None, f"rank {rank}:\n{err}"
