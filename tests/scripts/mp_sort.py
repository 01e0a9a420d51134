"""Multi-rank variable-length Sort job correctness script (torchrun, one process per GPU)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch
import torch.distributed as dist


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 40_000
    kind = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    side = (int(sys.argv[3]) if len(sys.argv) > 3 else 1) != 0
    lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    from uda_b200.models.sortjob import SortJob

    job = SortJob(n, 3, device=dev, key_kind=kind, key_len=(0, 60), val_len=(0, 120), tie_fraction=0.05, side_index=side).setup()
    job.step()
    out = job.step()
    res = job.check(out)
    res["rank"] = dist.get_rank()
    print("MPRESULT " + json.dumps(res), flush=True)
    ok = res["ok"]
    job.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
