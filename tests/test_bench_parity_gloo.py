"""bench.py's on-hardware value check (bench_workloads.parity_check) itself, on 4 CPU ranks against
tests/fake_peer.py: BASELINE configs 2-5 at the check's own shape (rows capped, batch 256 per rank).
It must come out ok, routing included, and it must notice a wrong table."""
import os
import uuid

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from test_sharded_gloo import free_port

W4 = 4


def run_rank(rank, port, tag, cfg, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=W4)
    mem = None
    try:
        import bench_workloads as bw
        from fake_peer import FakeMem, FakePeerOps
        mem = FakeMem(tag, rank, W4)
        bw.PARITY_ROW_CAP, bw.PARITY_B = 2000, 24           # the check's logic, not its size, is under test here
        res = bw.parity_check(cfg, W4, rank, torch.device("cpu"), kernel_ops=FakePeerOps(mem), peer_memory=True)
        assert res["ok"] and res["routing_bit_exact"] is True, res
        assert res["max_rel_err_fwd"] <= res["tolerance"] and res["max_rel_err_rows"] <= res["tolerance"]
        results[rank] = "ok"
    finally:
        dist.barrier()
        if mem is not None:
            mem.cleanup()
        dist.destroy_process_group()


@pytest.mark.parametrize("cfg", [2, 3, 4, 5])
def test_bench_parity_check_on_cpu_ranks(cfg):
    port = free_port()
    tag = "tdfo_fake_" + uuid.uuid4().hex[:8]
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank, args=(r, port, tag, cfg, results)) for r in range(W4)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=300)
        for p in procs:
            if p.is_alive():
                p.terminate()
        for f in os.listdir("/dev/shm"):
            if f.startswith(tag):
                os.remove(os.path.join("/dev/shm", f))
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {r: "ok" for r in range(W4)}
