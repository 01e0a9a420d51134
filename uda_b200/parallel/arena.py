"""Symmetric receive arena: one HBM region per rank, peer-mapped into every other rank.

Reference behaviour being replaced: the consumer registers one big memory region
(``(num_kv_bufs + 10) * 2 * buffer_size``, src/Merger/reducer.cc:453-496) and hands its address +
rkey to the provider in every fetch request (src/DataNet/RDMAClient.cc:559-600); the provider keeps
1000 registered chunks (src/DataNet/RDMAServer.cc:383-384). Here each rank allocates an arena with
``cudaMalloc`` (outside the torch caching allocator so it can be exported whole), exports a CUDA IPC
handle, all-gathers the handles over the torch process group (NCCL is used for this bootstrap only)
and opens every peer's arena. From then on a peer's arena is an ordinary device pointer: kernels store
into it over NVLink and publish flags that live in the same arena.

Arena layout (offsets in bytes)::

    0      barrier pad   64 x u64   slot[r] written by rank r (uda_barrier_all)
    512    status words  16 x u32, then the barrier epoch counter (u64 at 576) and this rank's count of merged
                         sub-partitions (u64 at 640)
    1024   credits       64 x u64   slot[r] = sub-partitions reducer r has merged so far (written by rank r into
                         every peer's arena: uda_credit_publish; read by the push kernel before it refills a ring slot)
    4096   flags         max_runs x u64  "bytes landed" counter per incoming run
    D      data          receive segments (D = 4096-aligned)
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
import torch.distributed as dist

from .._lib import check, kernels

BARRIER_OFF = 0
STATUS_OFF = 512
CREDIT_OFF = 1024
FLAGS_OFF = 4096


class RawDeviceBuffer:
    """A cudaMalloc'd (or IPC-imported) region exposed to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int, owner: bool, imported: bool = False):
        self.ptr, self.nbytes, self._owner, self._imported = ptr, nbytes, owner, imported
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None,
        }

    def tensor(self, device) -> torch.Tensor:
        return torch.as_tensor(self, device=device)

    def close(self) -> None:
        if self.ptr:
            if self._imported:
                kernels().uda_ipc_close_handle(self.ptr)
            elif self._owner:
                kernels().uda_dev_free(self.ptr)
            self.ptr = 0


def dev_malloc(nbytes: int) -> RawDeviceBuffer:
    """cudaMalloc outside the torch allocator. A failure is an :class:`ArenaAllocationFailed` (the counterpart of
    the reference's MR registration failure, src/DataNet/RDMAComm.cc:67-154), which the host answers with fallback."""
    from ..runtime.failure import ArenaAllocationFailed

    p = C.c_void_p()
    rc = kernels().uda_dev_malloc(C.byref(p), nbytes)
    if rc != 0 or not p.value:
        raise ArenaAllocationFailed(f"uda_dev_malloc({nbytes} bytes) failed: CUDA error {rc}")
    return RawDeviceBuffer(p.value, nbytes, owner=True)


class SymmetricArena:
    """``nbytes`` of HBM on every rank of ``group``, each mapped into all ranks."""

    def __init__(self, data_bytes: int, max_runs: int, device: torch.device, group: Optional[dist.ProcessGroup] = None):
        self.device = device
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.max_runs = max_runs
        self.data_off = (FLAGS_OFF + 8 * max_runs + 4095) // 4096 * 4096
        self.nbytes = self.data_off + (data_bytes + 4095) // 4096 * 4096
        self.local = dev_malloc(self.nbytes)
        check(kernels().uda_memset_async(self.local.ptr, 0, self.data_off, torch.cuda.current_stream().cuda_stream), "memset")
        torch.cuda.synchronize(device)
        self.peers: List[RawDeviceBuffer] = []
        self.peer_ptrs: List[int] = [0] * self.world
        self.peer_ptrs[self.rank] = self.local.ptr
        if self.world > 1:
            hb = kernels().uda_ipc_handle_bytes()
            mine = (C.c_ubyte * hb)()
            check(kernels().uda_ipc_get_handle(self.local.ptr, mine), "uda_ipc_get_handle")
            t = torch.tensor(list(mine), dtype=torch.uint8, device=device)
            allh = torch.empty(self.world * hb, dtype=torch.uint8, device=device)
            dist.all_gather_into_tensor(allh, t, group=group)
            allh = allh.cpu().numpy().reshape(self.world, hb)
            for r in range(self.world):
                if r == self.rank:
                    continue
                h = (C.c_ubyte * hb)(*allh[r].tolist())
                p = C.c_void_p()
                check(kernels().uda_ipc_open_handle(h, C.byref(p)), f"uda_ipc_open_handle(rank {r})")
                self.peers.append(RawDeviceBuffer(p.value, self.nbytes, owner=False, imported=True))
                self.peer_ptrs[r] = p.value
            dist.barrier(group=group)
        # device table of barrier-pad pointers (one per rank)
        self.pads_dev = torch.tensor([p + BARRIER_OFF for p in self.peer_ptrs], dtype=torch.int64, device=device)
        self.status = self.local_tensor()[STATUS_OFF:STATUS_OFF + 8].view(torch.int32)
        #: barrier epoch, kept on the device so that a barrier launch carries no host state (CUDA-graph replay)
        self.epoch_dev = self.local_tensor()[STATUS_OFF + 64:STATUS_OFF + 72].view(torch.int64)
        #: receive-ring flow control: credits[r] = sub-partitions reducer r has merged (r writes it here), merged_dev = my own count
        self.credits = self.local_tensor()[CREDIT_OFF:CREDIT_OFF + 8 * 64].view(torch.int64)
        self.merged_dev = self.local_tensor()[STATUS_OFF + 128:STATUS_OFF + 136].view(torch.int64)
        self.credit_slots_dev = torch.tensor([p + CREDIT_OFF + 8 * self.rank for p in self.peer_ptrs], dtype=torch.int64, device=device)

    # ---- views -------------------------------------------------------------------------------
    def local_tensor(self) -> torch.Tensor:
        return self.local.tensor(self.device)

    def flags(self) -> torch.Tensor:
        return self.local_tensor()[FLAGS_OFF:FLAGS_OFF + 8 * self.max_runs].view(torch.int64)

    def data(self) -> torch.Tensor:
        return self.local_tensor()[self.data_off:]

    def peer_data_ptr(self, r: int) -> int:
        return self.peer_ptrs[r] + self.data_off

    def peer_flag_ptr(self, r: int, run: int) -> int:
        return self.peer_ptrs[r] + FLAGS_OFF + 8 * run

    # ---- sync --------------------------------------------------------------------------------
    def reset_flags(self) -> None:
        self.flags().zero_()

    def barrier(self) -> None:
        """All-rank barrier over peer memory on the current stream (no NCCL, no host sync)."""
        from ..ops.kernels import barrier_all_counted

        barrier_all_counted(self.pads_dev, self.rank, self.world, self.epoch_dev, self.status)

    def publish_credit(self) -> None:
        """One more sub-partition of this reducer has been merged (its ring slot is free): tell every sender. Current stream."""
        from ..ops.kernels import credit_publish

        credit_publish(self.credit_slots_dev, self.world, self.merged_dev)

    def close(self) -> None:
        for p in self.peers:
            p.close()
        self.peers = []
        self.local.close()
