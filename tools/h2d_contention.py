"""Diagnostic: pinned host -> device copy bandwidth per rank when all ranks of the box copy at once
(the e2e arm of bench.py moves 68 MB of ids per step per GPU).
  torchrun --nproc-per-node 8 tools/h2d_contention.py"""
import os
import time

import torch
import torch.distributed as dist


def main():
    rank, local = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    n = 8_519_680
    host = torch.randint(0, 1_000_000, (n,), dtype=torch.int64).pin_memory()
    d = torch.empty(n, dtype=torch.int64, device=dev)
    for mode in ("all ranks at once", "one rank at a time"):
        res = []
        for turn in range(dist.get_world_size() if mode != "all ranks at once" else 1):
            dist.barrier()
            torch.cuda.synchronize()
            if mode == "all ranks at once" or turn == rank:
                for _ in range(3):
                    d.copy_(host, non_blocking=True)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(20):
                    d.copy_(host, non_blocking=True)
                e1.record()
                torch.cuda.synchronize()
                res.append(e0.elapsed_time(e1) / 20)
            dist.barrier()
        ms = res[0]
        out = [None] * dist.get_world_size()
        dist.all_gather_object(out, (rank, ms, n * 8 / ms / 1e6))
        if rank == 0:
            print(mode, " ".join(f"r{r}:{m:.2f}ms/{g:.1f}GB/s" for r, m, g in out), flush=True)
    dist.destroy_process_group()
    os._exit(0)


if __name__ == "__main__":
    main()
