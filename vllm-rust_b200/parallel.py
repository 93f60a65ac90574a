"""Tensor-parallel W4A16 linears: the composition the reference never executes
(its Column/RowParallelLinear hold dense weights, distributed/parallel_layers.rs:36-324)
built with the reference's own sharding and collective semantics:

  column: rank r keeps output columns [r*N/tp, (r+1)*N/tp), no collective
          (gather_output=false, distributed/attention.rs:76)
  row   : rank r keeps input rows [r*K/tp, (r+1)*K/tp) and the matching group rows;
          forward = local GEMM -> all_reduce(Sum) -> + bias   (parallel_layers.rs:275-313)

DeviceCommunicator mirrors distributed/communicator.rs; NcclCommunicator is the
library's dlopen'ed NCCL (nccl.rs:451-547, 1338-1364), bootstrapped from a 128-byte
unique id broadcast out of band (distributed/launcher.rs:204-262; here over
torch.distributed's store instead of raw TCP).
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from ._lib import W4A16Error, check, lib
from .quantization import TP_COLUMN, TP_ROW, B200W4A16Linear, QuantizedLinear


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """[start, end) of rank's contiguous shard (parallel_layers.rs:62-70, 205-213)."""
    if total % world:
        raise ValueError(f"dimension {total} is not divisible by world size {world}")
    per = total // world
    return rank * per, (rank + 1) * per


class DeviceCommunicator:
    """distributed/communicator.rs DeviceCommunicator (the part this path uses)."""

    def rank(self) -> int:
        raise NotImplementedError

    def world_size(self) -> int:
        raise NotImplementedError

    def all_reduce(self, t: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError

    def nccl_comm(self) -> Optional[int]:
        return None


class MockCommunicator(DeviceCommunicator):
    """communicator.rs:100-236: all_reduce is the identity."""

    def __init__(self, rank: int = 0, world_size: int = 1):
        self._r, self._w = rank, world_size

    def rank(self):
        return self._r

    def world_size(self):
        return self._w

    def all_reduce(self, t):
        return t


class TorchDistCommunicator(DeviceCommunicator):
    """all_reduce(Sum) through torch.distributed (gloo on CPU for the host-logic tests)."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self._dist, self._group = dist, group

    def rank(self):
        return self._dist.get_rank(self._group)

    def world_size(self):
        return self._dist.get_world_size(self._group)

    def all_reduce(self, t):
        self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self._group)
        return t


class NcclCommunicator(DeviceCommunicator):
    """ncclComm_t owned by libw4a16_b200.so; the id travels over torch.distributed."""

    def __init__(self, rank: int, world_size: int, device: int, unique_id: Optional[bytes] = None):
        self._r, self._w = rank, world_size
        self._comm = ctypes.c_void_p()
        if unique_id is None:
            import torch.distributed as dist
            buf = torch.zeros(128, dtype=torch.uint8)
            if rank == 0:
                raw = (ctypes.c_char * 128)()
                check(lib().w4a16_nccl_unique_id(ctypes.cast(raw, ctypes.c_void_p)))
                buf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
            if dist.get_backend() == "nccl":
                buf = buf.cuda(device)
            dist.broadcast(buf, src=0)
            unique_id = bytes(buf.cpu().tolist())
        raw = (ctypes.c_char * 128).from_buffer_copy(unique_id)
        with torch.cuda.device(device):
            check(lib().w4a16_nccl_comm_init(ctypes.byref(self._comm), world_size, ctypes.cast(raw, ctypes.c_void_p), rank))

    def rank(self):
        return self._r

    def world_size(self):
        return self._w

    def nccl_comm(self):
        return self._comm.value

    def all_reduce(self, t):
        raise W4A16Error("NcclCommunicator reduces inside w4a16_forward_row_allreduce; use RowParallelW4A16Linear")

    def destroy(self):
        if self._comm:
            lib().w4a16_nccl_comm_destroy(self._comm)
            self._comm = ctypes.c_void_p()


class PeerGroup:
    """Symmetric NVLink peer buffers of a tensor-parallel group (w4a16_peer_*): what the fused GEMV + all-reduce
    epilogue writes into.  Replaces the reference's dormant custom all-reduce (kernels/custom_allreduce.cu) for
    the decode sizes.  The 64-byte IPC handles travel over torch.distributed, like the NCCL unique id."""

    HANDLE_BYTES = 64

    def __init__(self, rank: int, world_size: int, device: int, max_out_features: int, group=None):
        """Collective: every rank of the group must call it.  A failure on ANY rank (no peer access, IPC refused)
        raises W4A16Error on EVERY rank after the exchange, so callers can fall back to NCCL consistently."""
        import torch.distributed as dist
        self._r, self._w = rank, world_size
        self._peer = ctypes.c_void_p()
        on_gpu = dist.get_backend(group) == "nccl"

        def gather(payload: bytes):
            mine = torch.frombuffer(bytearray(payload), dtype=torch.uint8).clone()
            if on_gpu:
                mine = mine.cuda(device)
            out = [torch.empty_like(mine) for _ in range(world_size)]
            dist.all_gather(out, mine, group=group)
            return [bytes(t.cpu().tolist()) for t in out]

        import contextlib
        with (torch.cuda.device(device) if torch.cuda.is_available() else contextlib.nullcontext()):
            err = ""
            raw = (ctypes.c_char * self.HANDLE_BYTES)()
            try:
                check(lib().w4a16_peer_create(ctypes.byref(self._peer), rank, world_size, int(max_out_features) * 4))
                check(lib().w4a16_peer_export(self._peer, ctypes.cast(raw, ctypes.c_void_p)))
            except W4A16Error as e:
                err = str(e) or "peer_create failed"
            got = gather(bytes(raw.raw) + (b"\x01" if err else b"\x00"))
            if not err and not any(g[-1] for g in got):
                blob = b"".join(g[:self.HANDLE_BYTES] for g in got)
                buf = (ctypes.c_char * len(blob)).from_buffer_copy(blob)
                try:
                    check(lib().w4a16_peer_attach(self._peer, ctypes.cast(buf, ctypes.c_void_p)))
                except W4A16Error as e:
                    err = str(e) or "peer_attach failed"
            elif not err:
                err = "a peer rank could not create its buffer"
            # second exchange doubles as the barrier: nobody launches before every rank has mapped every buffer
            bad = [i for i, g in enumerate(gather(b"\x01" if err else b"\x00")) if g[0]]
            if bad:
                self.destroy()
                raise W4A16Error(f"PeerGroup: ranks {bad} failed ({err or 'see those ranks'})")

    def ptr(self) -> Optional[int]:
        return self._peer.value

    def destroy(self):
        if self._peer:
            lib().w4a16_peer_destroy(self._peer)
            self._peer = ctypes.c_void_p()


class ColumnParallelW4A16Linear(QuantizedLinear):
    """ColumnParallelLinear (parallel_layers.rs:36-170) over a W4A16 shard; gather_output=false."""

    def __init__(self, config, in_features: int, out_features: int, bias: bool, comm: DeviceCommunicator, device):
        self.comm = comm
        self.inner: B200W4A16Linear = config.create_linear(
            in_features, out_features, bias, device, tp_mode=TP_COLUMN, tp_rank=comm.rank(), tp_size=comm.world_size())

    def load_weights(self, weights):
        self.inner.load_weights(weights)  # FULL checkpoint tensors; the shard reads its slice

    def forward(self, x):
        return self.inner.forward(x)

    def forward_into(self, x2, y):
        self.inner.forward_into(x2, y)

    def in_features(self):
        return self.inner.in_features()

    def out_features(self):
        return self.inner.out_features()

    def has_bias(self):
        return self.inner.has_bias()

    def weight_dtype(self):
        return self.inner.weight_dtype()

    def kernel_for(self, m: int) -> str:
        return self.inner.kernel_for(m)


class RowParallelW4A16Linear(QuantizedLinear):
    """RowParallelLinear (parallel_layers.rs:181-324): local GEMM, all_reduce(Sum), + bias."""

    def __init__(self, config, in_features: int, out_features: int, bias: bool, comm: DeviceCommunicator, device,
                 peers: Optional["PeerGroup"] = None):
        self.comm = comm
        self.peers = peers   # with a PeerGroup decode-sized calls all-reduce inside the GEMV kernel (no NCCL launch)
        self.inner: B200W4A16Linear = config.create_linear(
            in_features, out_features, bias, device, tp_mode=TP_ROW, tp_rank=comm.rank(), tp_size=comm.world_size())

    def load_weights(self, weights):
        self.inner.load_weights(weights)

    def forward(self, x):
        self.inner._check_x(x)
        lead = x.shape[:-1]
        x2 = x.contiguous().view(-1, x.shape[-1])
        y = torch.empty((x2.shape[0], self.out_features()), dtype=torch.bfloat16, device=x.device)
        self.forward_into(x2, y)
        return y.view(*lead, self.out_features())

    def forward_into(self, x2, y):
        if self.comm.world_size() > 1 and self.comm.nccl_comm() is None:
            raise W4A16Error("RowParallelW4A16Linear needs an NcclCommunicator for world_size > 1")
        if self.peers is not None and self.comm.world_size() > 1:
            self.inner.forward_row_fused(x2, y, self.peers.ptr(), self.comm.nccl_comm())
        else:
            self.inner.forward_row_allreduce(x2, y, self.comm.nccl_comm())

    def in_features(self):
        return self.inner.in_features()

    def out_features(self):
        return self.inner.out_features()

    def has_bias(self):
        return self.inner.has_bias()

    def weight_dtype(self):
        return self.inner.weight_dtype()

    def kernel_for(self, m: int) -> str:
        return self.inner.kernel_for(m)
