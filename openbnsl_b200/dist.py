"""torch.distributed plumbing for the multi-GPU drivers.

The C++ drivers shard a batch of tuples over ranks and need exactly two small
collectives: all-reduce(min) of int32 "first passing
candidate per edge" for PC, and an all-gather of score rows per lattice level
for exhaustive search.  ``TorchComm`` adapts ``torch.distributed`` (NCCL over
NVLink on GPUs, gloo in the CPU test-suite) to the ``comm`` argument of
``openbnsllib.structure_learning.pc`` / ``exhaustive_search`` -- this is the HOST-SIDE communicator
used by the CPU test-suite (gloo, world size 2) and kept as an A/B; the product path on GPUs is
``init_native_comm`` below: NCCL owned by the library, below the C boundary.
"""
from __future__ import annotations

import numpy as np


def init_native_comm(lib=None, group=None):
    """Give this process's ctx the library's OWN NCCL communicator (obn_comm_init) and return the
    ``comm=`` argument that selects it.  torch.distributed is used for one thing only: carrying the
    128-byte NCCL unique id from rank 0 to the other ranks (any host-side channel would do).  After
    this call the data-path collectives -- all-reduce(min) per PC round, all-gathers of the
    exhaustive search -- are issued by libopenbnsl_b200 on its own device buffers and stream."""
    import torch.distributed as dist
    if lib is None:
        import openbnsl_b200
        lib = openbnsl_b200.openbnsllib
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    box = [lib.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    lib.comm_init(box[0], rank, world)
    return lib.nccl_comm()


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
