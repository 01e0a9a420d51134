"""Multi-rank TeraSort correctness script (launched by torchrun, one process per GPU)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch
import torch.distributed as dist


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    maps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    subparts = int(sys.argv[4]) if len(sys.argv) > 4 else 1
    mode = sys.argv[5] if len(sys.argv) > 5 else "lsu"
    zero_copy = len(sys.argv) > 6 and sys.argv[6] == "1"
    slots = int(sys.argv[7]) if len(sys.argv) > 7 else 0          # receive slots (< subparts: ring with credits)
    local_copy = sys.argv[8] if len(sys.argv) > 8 else "auto"
    lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    from uda_b200.models.terasort import TeraSortJob

    job = TeraSortJob(n, maps, device=dev, subparts=subparts, push_mode=mode, push_ctas=16 if mode == "tma" else 0, local_zero_copy=zero_copy, arena_slots=slots,
                      local_copy=local_copy).setup()
    out = None
    for _ in range(steps):
        out = job.step()
    res = job.check(out)
    torch.cuda.synchronize()
    # P2P sanity: arena peer pointers differ and the barrier works
    job.arena.barrier()
    torch.cuda.synchronize()
    res["barrier_status"] = int(job.arena.status[0].item())
    res["rank"] = dist.get_rank()
    res["n_out"] = job.n_out
    res["arena_bytes"] = int(job.plan.max_arena_bytes())
    res["slots"] = int(job.plan.slots)
    print("MPRESULT " + json.dumps(res), flush=True)
    ok = res["ok"] and res["barrier_status"] == 0
    job.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
