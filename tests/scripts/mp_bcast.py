"""Side-table broadcast check (torchrun): NVLS multicast when available, unicast peer stores otherwise."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch
import torch.distributed as dist


def main():
    lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    from uda_b200.parallel.nvls import TableBroadcaster

    res = {"rank": dist.get_rank()}
    ok = True
    for allow in (True, False):
        nbytes = 40_000
        b = TableBroadcaster(nbytes, dev, allow_nvls=allow)
        for it, root in enumerate((0, dist.get_world_size() - 1)):
            g = torch.Generator(device="cpu").manual_seed(100 + it)
            table = torch.randint(0, 256, (nbytes - 7,), dtype=torch.uint8, generator=g)
            got = b.broadcast(table.to(dev) if dist.get_rank() == root else None, root=root)
            torch.cuda.synchronize()
            ok = ok and bool(torch.equal(got[: nbytes - 7].cpu(), table))
        res["mode_allow_nvls" if allow else "mode_forced_unicast"] = b.mode
        if allow and b.why:
            res["nvls_unavailable_because"] = b.why[:200]
        b.close()
    res["ok"] = ok
    print("MPRESULT " + json.dumps(res), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
