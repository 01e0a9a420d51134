"""Broadcast of small side tables (partition splitters, bucket-boundary keys, configuration) to every rank.

No counterpart in the reference (Hadoop ships such tables as distributed-cache files). On NVSwitch systems the
table is written once with ``multimem.st`` to a multicast address, and the switch replicates it into every rank's
copy of a symmetric buffer (NVLS). The multicast mapping comes from torch's symmetric-memory rendezvous; when the
platform offers none (no NVSwitch, multicast disabled, or the rendezvous fails inside a container) the same call
falls back to unicast peer stores into the IPC arena. ``mode`` tells which path is in use.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from .._lib import check, kernels
from .arena import SymmetricArena


class TableBroadcaster:
    def __init__(self, nbytes: int, device: torch.device, group=None, allow_nvls: bool = True):
        self.device, self.group = device, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.nbytes = (nbytes + 15) // 16 * 16
        self.arena = SymmetricArena(self.nbytes, 1, device, group)   # barrier pads + unicast landing buffer
        self.peers_dev = torch.tensor([self.arena.peer_data_ptr(p) for p in range(self.world)], dtype=torch.int64, device=device)
        self.mode, self.why = "unicast", ""
        self._symm, self._hdl, self._mc = None, None, 0
        if allow_nvls and self.world > 1:
            try:
                import torch.distributed._symmetric_memory as symm_mem

                t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=device)
                hdl = symm_mem.rendezvous(t, group=dist.group.WORLD if group is None else group)
                mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
                if mc:
                    self._symm, self._hdl, self._mc, self.mode = t, hdl, mc, "nvls"
                else:
                    self.why = "symmetric memory has no multicast mapping on this platform"
            except Exception as e:  # keep working without NVLS, but say why
                self.why = f"{type(e).__name__}: {e}"
        # every rank must agree on the path
        if self.world > 1:
            flag = torch.tensor([1 if self.mode == "nvls" else 0], dtype=torch.int32, device=device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
            if int(flag.item()) == 0 and self.mode == "nvls":
                self.mode, self.why = "unicast", "another rank has no multicast mapping"

    def local(self) -> torch.Tensor:
        """This rank's copy of the table."""
        return self._symm if self.mode == "nvls" else self.arena.data()[: self.nbytes]

    def broadcast(self, table: Optional[torch.Tensor], root: int = 0) -> torch.Tensor:
        """``table`` (uint8, <= nbytes, on the root) appears in every rank's :meth:`local` buffer. Stream ordered on
        the current stream; the all-rank barrier inside makes the data visible before it returns control to the stream."""
        stream = torch.cuda.current_stream(self.device).cuda_stream
        if self.world > 1:
            self.arena.barrier()                                  # previous readers are done with the buffer
        if self.rank == root:
            staged = torch.zeros(self.nbytes, dtype=torch.uint8, device=self.device)
            staged[: table.numel()].copy_(table.reshape(-1).view(torch.uint8))
            if self.mode == "nvls":
                check(kernels().uda_multicast_store(staged.data_ptr(), self._mc, self.nbytes, stream), "uda_multicast_store")
            else:
                check(kernels().uda_unicast_store(staged.data_ptr(), self.peers_dev.data_ptr(), self.world, self.nbytes, stream),
                      "uda_unicast_store")
        if self.world > 1:
            self.arena.barrier()                                  # release/acquire: the table has landed everywhere
        return self.local()

    def close(self) -> None:
        self.arena.close()
