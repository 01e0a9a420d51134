#!/usr/bin/env python
"""TeraSort shuffle + merge on one B200, step by step (needs a GPU; `torchrun --nproc-per-node N` for several).

    python examples/terasort_one_gpu.py [records] [maps]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from uda_b200.models.terasort import TeraSortJob
from uda_b200.runtime.failure import FallbackController
from uda_b200.utils.trace import PhaseTimer


def main() -> int:
    records = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
    maps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    dev = torch.device("cuda", 0)
    fallback = FallbackController(on_failure=lambda e: print("would fall back to the vanilla shuffle:", e))
    with fallback.guard("init"):
        job = TeraSortJob(records, maps, device=dev).setup()     # teragen -> map-side sort -> MOFs in HBM -> arena
    if fallback.failed:
        return 1
    timer = PhaseTimer(dev)
    with fallback.guard("shuffle+merge"):
        for _ in range(3):
            with timer.phase("step", bytes_moved=job.record_bytes_local):
                merged = job.step()                               # push -> landed-byte acquire -> record merge
        job.raise_on_failure()                                    # watchdog time-outs surface here
        job.verify_crc()                                          # landed segments == what the maps wrote
    print("validation:", job.check(merged))
    print("timing:", timer.summary())
    print("first record key:", bytes(merged[2:12].cpu().numpy()).hex(), " stream bytes:", merged.numel())
    job.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
