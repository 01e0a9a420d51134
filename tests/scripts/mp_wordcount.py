"""Multi-rank Zipf WordCount correctness script (torchrun, one process per GPU)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch
import torch.distributed as dist


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    vocab = int(sys.argv[2]) if len(sys.argv) > 2 else 20_000
    lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    from uda_b200.models.wordcount import WordCountJob

    job = WordCountJob(n, vocab=vocab, device=dev).setup()
    job.step()
    job.step()
    res = job.check()
    res["rank"] = dist.get_rank()
    print("MPRESULT " + json.dumps(res), flush=True)
    ok = res["ok"]
    job.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
