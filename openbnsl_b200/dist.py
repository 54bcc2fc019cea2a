"""torch.distributed plumbing for the multi-GPU drivers.

The C++ drivers shard a batch of tuples over ranks and need exactly two small
collectives: all-reduce(min) of int32 "first passing
candidate per edge" for PC, and an all-gather of score rows per lattice level
for exhaustive search.  ``TorchComm`` adapts ``torch.distributed`` (NCCL over
NVLink on GPUs, gloo in the CPU test-suite) to the ``comm`` argument of
``openbnsllib.structure_learning.pc`` / ``exhaustive_search``.
"""
from __future__ import annotations

import numpy as np


class TorchComm:
    def __init__(self, group=None, device=None):
        import torch
        import torch.distributed as dist
        self._torch, self._dist = torch, dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        backend = dist.get_backend(group)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
        self.device = device
        self.n_collectives = 0

    def allreduce_min_i32(self, arr: np.ndarray) -> None:
        torch, dist = self._torch, self._dist
        t = torch.from_numpy(arr)
        if self.device.type == "cuda":
            d = t.to(self.device, non_blocking=False)
            dist.all_reduce(d, op=dist.ReduceOp.MIN, group=self.group)
            t.copy_(d)
        else:
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        self.n_collectives += 1

    def allgather(self, send: np.ndarray, recv: np.ndarray) -> None:
        torch, dist = self._torch, self._dist
        s, r = torch.from_numpy(send), torch.from_numpy(recv)
        if self.device.type == "cuda":
            sd = s.to(self.device)
            rd = torch.empty(r.shape, dtype=r.dtype, device=self.device)
            dist.all_gather_into_tensor(rd, sd, group=self.group)
            r.copy_(rd)
        else:
            dist.all_gather_into_tensor(r, s, group=self.group)
        self.n_collectives += 1

    def allgather_dev(self, ptr: int, bytes_per_rank: int) -> None:
        """In-place all-gather of a device buffer of world * bytes_per_rank bytes (rank r owns
        slice r).  The buffer belongs to libopenbnsl_b200; it is wrapped zero-copy through
        __cuda_array_interface__ and handed to NCCL."""
        torch, dist = self._torch, self._dist
        if self.device.type != "cuda":
            raise RuntimeError("allgather_dev needs the nccl backend")

        class _Buf:
            pass
        b = _Buf()
        b.__cuda_array_interface__ = {"shape": (self.world * bytes_per_rank,), "typestr": "|u1",
                                      "data": (int(ptr), False), "version": 2}
        full = torch.as_tensor(b, device=self.device)
        mine = full[self.rank * bytes_per_rank:(self.rank + 1) * bytes_per_rank]
        torch.cuda.current_stream().synchronize()
        dist.all_gather_into_tensor(full, mine, group=self.group)
        torch.cuda.current_stream().synchronize()
        self.n_collectives += 1
