"""Real ranks on real GPUs: a tensor-parallel decoder-layer chain in ONE launch per rank (w4a16_chain_set_peer):
[qkv column shard -> o row shard (all-reduce inside the launch) -> gate_up column shard -> down row shard (all-reduce)],
twice, every linear reading what the previous one wrote.  Checked on every rank against the float64 oracle of the FULL
(unsharded) linears applied to the chain's own intermediate buffers, bit-identity of the all-reduced outputs across
ranks, repeated runs on changing inputs (buffer rotation), CUDA-graph replay, and a shard width that is not a multiple
of 4096.  Skipped on a box with fewer GPUs than ranks."""
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
        g, H, I, QKV = 128, 4096, 8192 + 1024 * world, 4096 + 2048      # I / world is not a multiple of 4096 at world 2
        comm = P.NcclCommunicator(rank, world, rank)
        peers = P.PeerGroup(rank, world, rank, H)
        cfg = P.AwqConfig.int4(g)
        # (name, K, N, mode, bias, weight scale): small scales keep the magnitudes of the eight-linear chain in range
        spec = [("qkv", H, QKV, "col", False), ("o", H, H, "row", True), ("gate_up", H, I, "col", False), ("down", I, H, "row", False)]
        lins, cks = [], []
        for layer in range(2):
            for i, (name, K, N, mode, bias) in enumerate(spec):
                qweight, qzeros, scales, b = O.make_awq_checkpoint(K, N, g, seed=400 + 10 * layer + i, bias=bias)   # same on every rank
                scales = (scales * np.float16(0.2)).astype(np.float16)
                w = {"qweight": t_i32(qweight).to(dev), "qzeros": t_i32(qzeros).to(dev), "scales": torch.from_numpy(scales).to(dev)}
                if bias:
                    w["bias"] = torch.from_numpy(b).to(torch.bfloat16).to(dev)
                cls = P.ColumnParallelW4A16Linear if mode == "col" else P.RowParallelW4A16Linear
                lin = cls(cfg, K, N, bias, comm, dev)
                lin.load_weights(w)
                lins.append((name, K, N, mode, lin))
                cks.append(awq_unpacked(qweight, qzeros, scales) + (O.np_bf16_round(b) if bias else None,))
        M = 1
        x0 = torch.zeros((M, H), dtype=torch.bfloat16, device=dev)
        bufs, steps = [x0], []
        for name, K, N, mode, lin in lins:
            xin = bufs[-1][:, :lin.in_features()]        # a row shard reads the leading part of the column shard before it
            y = torch.zeros((M, lin.out_features()), dtype=torch.bfloat16, device=dev)
            steps.append((lin, xin, y))
            bufs.append(y)
        chain = P.DecodeChain(steps, peers=peers)
        name = chain.kernel()
        assert name.startswith("chain_imma") and "4 fused all-reduces" in name, name

        def check(tag):
            torch.cuda.synchronize()
            for i, (nm, K, N, mode, lin) in enumerate(lins):
                q, z, sb, bias = cks[i]
                xin = steps[i][1]
                got = steps[i][2]
                n0, n1 = (P.shard_range(N, rank, world) if mode == "col" else (0, N))
                if mode == "col":
                    want = oracle_forward(bits_of(xin), q[:, n0:n1], z[:, n0:n1], sb[:, n0:n1], g)
                else:
                    # the full input of a row-parallel linear = the concatenation of every rank's input slice
                    parts = [torch.empty_like(xin) for _ in range(world)]
                    dist.all_gather(parts, xin.contiguous())
                    full = torch.cat(parts, dim=1)
                    want = oracle_forward(bits_of(full), q, z, sb, g, bias_f32=bias)
                    ys = [torch.empty_like(got) for _ in range(world)]
                    dist.all_gather(ys, got)
                    for other in ys:
                        assert torch.equal(other.view(torch.int16), got.view(torch.int16)), f"{tag}: {nm} differs across ranks"
                assert float(np.abs(want).max()) > 0
                assert_close_ref(bits_of(got), want, f"{tag}: linear {i} ({nm}) rank {rank}")

        for it in range(5):                            # more runs than exchange buffers
            x0.copy_(t_bf16_from_bits(O.make_activations(M, H, seed=500 + it)).to(dev))
            chain.run()
            check(f"run {it}")
        st = torch.cuda.Stream(dev)
        with torch.cuda.stream(st):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=st):
                chain.run()
                chain.run()
            x0.copy_(t_bf16_from_bits(O.make_activations(M, H, seed=600)).to(dev))
            graph.replay()
            graph.replay()
        st.synchronize()
        check("graph replay")
        dist.barrier()
        del chain
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
def test_tp_chain_real_ranks(world):
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    try:
        results = [q.get(timeout=300) for _ in range(world)]
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.terminate()
    for rank, err in results:
        assert err is None, f"rank {rank}:\n{err}"
