"""The peer-memory branch of the sharded module at world sizes the 2-GPU tests cannot reach.

On the GPUs the table- / column-wise exchange is done by kernels storing into NVLink peer memory at
addresses the host computes (`_peer_meta`, `_peer_grad_meta`, landing-zone offsets, rotating output
buffers).  tests/test_gpu_sharded.py checks that with W = 2; bench.py runs it with W = 4 and 8 but
checks no values.  Here the same host code runs on 4 CPU ranks against tests/fake_peer.py (files
under /dev/shm as "peer memory", the five peer kernels restated from their C-ABI contracts), with
uneven table-wise ownership, a table split into 4 column shards, a 2-shard one and a row-wise one.
Two training steps: outputs equal the unsharded oracle, tables equal the unsharded update."""
import os
import uuid

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle as orc
from test_sharded_gloo import free_port

B4, D4 = 5, 16
CASES = {
    # uneven table-wise ownership, two tables in 4 column shards each, one row-wise
    "w4": dict(W=4, rows=[23, 40, 9, 31, 17, 50, 12, 28]),
    # the shape bench.py runs on 8 GPUs: 26 equal-sized tables under balanced_plan (24 table-wise round-robin +
    # 2 tables x 4 column shards, one shard per rank)
    "w8_bench_plan": dict(W=8, rows=[11 + (7 * i) % 13 for i in range(26)]),
    # BASELINE config 4's plan: every table row-wise (partial sums land in per-source slots of the sample owner,
    # gradient rows are written into every rank's landing buffer)
    "w4_all_rw": dict(W=4, rows=[37, 50, 9, 64, 21]),
    # BASELINE config 5's plan on 8 ranks: 8 tables table-wise (one per rank) + 8 row-wise
    "w8_mixed": dict(W=8, rows=[20 + (5 * i) % 17 for i in range(8)] + [40 + (11 * i) % 23 for i in range(8)]),
}


def batch(rank, step, ROWS):
    rng = np.random.default_rng(1000 * step + rank)
    F = len(ROWS)
    lengths = rng.integers(0, 5, size=F * B4).astype(np.int32)
    lengths[(rank + step)::6] = 0
    off = orc.complete_cumsum(lengths)
    idx = np.concatenate([rng.integers(0, ROWS[f], size=int(off[(f + 1) * B4] - off[f * B4])) for f in range(F)])
    return lengths, idx.astype(np.int64), rng.normal(size=(B4, F * D4)).astype(np.float32)


def run_rank(rank, port, tag, case, results, prefetch=False, cuda=False):
    """cuda=True (tests/test_gpu_sharded.py): the same steps on W real GPUs, CUDA kernels over symmetric memory."""
    W4, ROWS = CASES[case]["W"], CASES[case]["rows"]
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    if cuda:
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=W4, device_id=dev)
    else:
        dev = torch.device("cpu")
        dist.init_process_group("gloo", rank=rank, world_size=W4)
    mem = None
    try:
        import tdfo_b200 as t
        from tdfo_b200 import distributed as td
        kernel_ops = None
        if not cuda:
            from fake_peer import FakeMem, FakePeerOps
            mem = FakeMem(tag, rank, W4)
            kernel_ops = FakePeerOps(mem)
        F = len(ROWS)
        tables = [t.EmbeddingBagConfig(num_embeddings=r, embedding_dim=D4, name=f"t{i}", feature_names=[f"f{i}"],
                                       pooling=t.PoolingType.SUM) for i, r in enumerate(ROWS)]
        if case == "w4":
            plan = {"t0": td.table_wise(2), "t1": td.column_wise([3, 1, 0, 2]), "t2": td.table_wise(0),
                    "t3": td.row_wise(), "t4": td.table_wise(2), "t5": td.column_wise([1, 2, 3, 0]),
                    "t6": td.table_wise(3), "t7": td.table_wise(2)}
        elif case == "w4_all_rw":
            plan = {c.name: td.row_wise() for c in tables}
        elif case == "w8_mixed":
            plan = {c.name: (td.table_wise(i) if i < 8 else td.row_wise()) for i, c in enumerate(tables)}
        else:
            plan = td.balanced_plan(tables, W4, min_shard_bytes=16)
        m = td.ShardedEmbeddingBagCollection(tables, plan, dev, kernel_ops=kernel_ops, init_seed=21, peer_memory=True)
        assert m._use_peer
        if case == "w4":
            assert m._n_tw == [1, 0, 3, 1] and m._n_cw == [2, 2, 2, 2] and m._Dc == 4
        elif case == "w4_all_rw":
            assert m._n_tw == [0] * 4 and m._n_rw == 5
        elif case == "w8_mixed":
            assert m._n_tw == [1] * 8 and m._n_rw == 8
        else:
            assert m._n_tw == [3] * 8 and m._n_cw == [1] * 8 and m._Dc == 4
        m.set_fused_optimizer(t.EmbOptimType.EXACT_SGD, learning_rate=0.05)
        m.train()
        w_ref = np.concatenate([td.ShardedEmbeddingBagCollection.initial_table(c, 21).numpy() for c in tables])
        base = np.concatenate([[0], np.cumsum(ROWS)[:-1]]).astype(np.int64)
        col = (np.arange(F) * D4).astype(np.int64)
        keys = [f"f{i}" for i in range(F)]
        mk = lambda b_: t.KeyedJaggedTensor.from_lengths_sync(keys, torch.from_numpy(b_[1]),
                                                              torch.from_numpy(b_[0])).to(dev)
        n_steps = 4 if prefetch else 2
        nxt = None
        for step in range(n_steps):
            lengths, idx, grad = batch(rank, step, ROWS)
            kjt = nxt if nxt is not None else mk((lengths, idx))
            out = m(kjt).values()
            nxt = None
            if prefetch and step + 1 < n_steps:
                # the next batch's exchange is issued before this batch's backward (double-buffered landing zones)
                nxt = mk(batch(rank, step + 1, ROWS))
                m.prefetch_input_dist(nxt)
                assert len(m._peer.zones) == 2
            want, _ = orc.embedding_forward(w_ref, D4, idx, orc.complete_cumsum(lengths), None, base, ROWS, col, F, B4,
                                            orc.POOL_SUM)
            np.testing.assert_allclose(out.detach().cpu().numpy(), want, rtol=1e-5, atol=1e-6, err_msg=f"step {step}")
            out.backward(torch.from_numpy(grad).to(dev))
            m.check_errors()
            # unsharded SGD on the global batch: per feature, rank 0's bags first
            per = [batch(r, step, ROWS) for r in range(W4)]
            offs = [orc.complete_cumsum(p[0]) for p in per]
            g_len = np.concatenate([per[r][0].reshape(F, B4)[f] for f in range(F) for r in range(W4)]).astype(np.int32)
            g_idx = np.concatenate([per[r][1][offs[r][f * B4]:offs[r][(f + 1) * B4]] for f in range(F) for r in range(W4)])
            g_grad = np.concatenate([p[2] for p in per], axis=0)
            orc.embedding_backward_update(w_ref, D4, None, None, g_grad, g_idx, orc.complete_cumsum(g_len), base, ROWS, col,
                                          F, W4 * B4, orc.POOL_SUM, orc.OptArgs(orc.OPT_SGD, 0.05))
        full = m.full_state_dict()
        for i, r in enumerate(ROWS):
            np.testing.assert_allclose(full[f"embedding_bags.t{i}.weight"].numpy(), w_ref[base[i]:base[i] + r],
                                       rtol=1e-5, atol=1e-6, err_msg=f"table t{i}")
        results[rank] = "ok"
    finally:
        dist.barrier()
        if mem is not None:
            mem.cleanup()
        dist.destroy_process_group()


@pytest.mark.parametrize("case,prefetch", [("w4", False), ("w8_bench_plan", False), ("w4", True),
                                           ("w4_all_rw", False), ("w8_mixed", False), ("w4_all_rw", True)])
def test_peer_memory_layout_beyond_two_ranks(case, prefetch):
    """prefetch=True: every step's input dist is issued through prefetch_input_dist before the previous
    backward, alternating between two landing zones."""
    W4 = CASES[case]["W"]
    port = free_port()
    tag = "tdfo_fake_" + uuid.uuid4().hex[:8]
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank, args=(r, port, tag, case, results, prefetch)) for r in range(W4)]
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
