"""Sharded checkpoints (`torch.save(model.state_dict(), ...)` on every rank at
torchrec/train.py:172-176), world_size 2 over gloo with the oracle-backed kernel table:

* state_dict() has the unsharded module's per-table keys, holding this rank's rows + shard metadata;
* full_state_dict() reassembles every table, equal to the tables the ranks were initialised from
  after training steps are applied to both;
* a module with ANOTHER plan loads either the full tables or its own shards and continues with
  identical outputs (resharding), and the fused optimizer state survives a save / load.
"""
import io
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from test_sharded_gloo import B, D, W, free_port, local_batch, make_tables


def _step(t, m, rank, rows):
    lengths, idx, grad = local_batch(rank, rows)
    kjt = t.KeyedJaggedTensor.from_lengths_sync([f"f{i}" for i in range(len(rows))], torch.from_numpy(idx),
                                                torch.from_numpy(lengths))
    out = m(kjt).values()
    out.backward(torch.from_numpy(grad))
    return out.detach().numpy().copy()


def run_rank(rank, port, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=W)
    try:
        import tdfo_b200 as t
        from tdfo_b200 import distributed as td
        from oracle_ops import OracleKernelOps
        tables = make_tables(t)
        rows = [c.num_embeddings for c in tables]
        plan_a = {"t0": td.table_wise(1), "t1": td.row_wise(), "t2": td.table_wise(0), "t3": td.row_wise(),
                  "t4": td.table_wise(1)}
        plan_b = {"t0": td.row_wise(), "t1": td.table_wise(0), "t2": td.row_wise(), "t3": td.table_wise(1),
                  "t4": td.column_wise([1, 0])}
        mk = lambda plan, seed: td.ShardedEmbeddingBagCollection(tables, plan, torch.device("cpu"),
                                                                 kernel_ops=OracleKernelOps(), init_seed=seed)
        a = mk(plan_a, 3)
        a.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=0.05)
        a.train()
        _step(t, a, rank, rows)

        # ---- per-rank state_dict: unsharded key names, local rows, metadata; survives torch.save
        sd = a.state_dict()
        assert [k for k in sd if k.endswith(".weight")] == [f"embedding_bags.t{i}.weight" for i in range(5)]
        meta = sd["_shard_metadata"]
        assert meta["world"] == W and meta["rank"] == rank
        assert sd["embedding_bags.t0.weight"].shape == ((50, D) if rank == 1 else (0, D))
        assert sd["embedding_bags.t1.weight"].shape == ((21, D) if rank == 0 else (20, D))
        assert meta["tables"]["t1"] == {"row_offset": 21 * rank, "rows": 21 - rank, "num_embeddings": 41}
        buf = io.BytesIO()
        torch.save({"model": sd, "optim": a.fused_optimizer.state_dict()}, buf)
        buf.seek(0)
        ck = torch.load(buf, weights_only=False)

        # ---- full tables, identical on both ranks
        full = a.full_state_dict()
        assert list(full) == [f"embedding_bags.t{i}.weight" for i in range(5)]
        for i, r_ in enumerate(rows):
            assert full[f"embedding_bags.t{i}.weight"].shape == (r_, D)
        gathered = [None] * W
        dist.all_gather_object(gathered, {k: v.numpy() for k, v in full.items()})
        for k in full:
            np.testing.assert_array_equal(gathered[0][k], gathered[1][k])
        for tname, info in meta["tables"].items():              # my shard is the slice the metadata names
            lo = info["row_offset"]
            np.testing.assert_array_equal(full[f"embedding_bags.{tname}.weight"][lo:lo + info["rows"]].numpy(),
                                          sd[f"embedding_bags.{tname}.weight"].numpy())

        # ---- same plan, fresh module: local shards + optimizer state restore the training state
        a2 = mk(plan_a, 99)
        a2.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=0.05)
        a2.train()
        a2.load_state_dict(ck["model"])
        a2.fused_optimizer.load_state_dict(ck["optim"])
        assert a2.fused_optimizer.iteration == a.fused_optimizer.iteration
        out_a, out_a2 = _step(t, a, rank, rows), _step(t, a2, rank, rows)
        np.testing.assert_array_equal(out_a, out_a2)
        fa, fa2 = a.full_state_dict(), a2.full_state_dict()       # the Adagrad state came along: same update
        for k in fa:
            np.testing.assert_array_equal(fa[k].numpy(), fa2[k].numpy())

        # ---- another plan loads the full tables (resharding) and an unsharded module's state_dict
        b = mk(plan_b, 77)
        b.load_state_dict(fa)
        b.eval()
        a.eval()
        lengths, idx, _ = local_batch(rank, rows)
        kjt = t.KeyedJaggedTensor.from_lengths_sync([f"f{i}" for i in range(5)], torch.from_numpy(idx),
                                                    torch.from_numpy(lengths))
        np.testing.assert_allclose(b(kjt).values().numpy(), a(kjt).values().numpy(), rtol=1e-6, atol=1e-6)
        sd_b = b.state_dict()                                     # a column shard: half the columns of every row
        assert sd_b["embedding_bags.t4.weight"].shape == (9, D // 2)
        assert sd_b["_shard_metadata"]["tables"]["t4"]["col_offset"] == (D // 2) * (1 - rank)
        b.load_state_dict(sd_b)                                   # own shards load back
        fb = b.full_state_dict()
        for k in fa:
            np.testing.assert_array_equal(fa[k].numpy(), fb[k].numpy())
        # the reference saves model.state_dict() to ONE path from every rank (torchrec/train.py:172-176), so
        # the last writer wins: the other rank's shard file must be refused, not silently loaded.  Row blocks
        # of t0 / t2 have the same shape on both ranks, so only the metadata can tell.
        theirs = [None] * W
        dist.all_gather_object(theirs, {k: (v.numpy() if torch.is_tensor(v) else v) for k, v in sd_b.items()})
        other = {k: (torch.from_numpy(v) if isinstance(v, np.ndarray) else v) for k, v in theirs[1 - rank].items()}
        try:
            b.load_state_dict(other)
            raise AssertionError("another rank's shards accepted")
        except RuntimeError as e:
            assert "shard mismatch" in str(e)
        try:
            b.load_state_dict({k: v[:3] for k, v in fa.items()})
            raise AssertionError("wrong-shaped checkpoint accepted")
        except RuntimeError as e:
            assert "size mismatch" in str(e)
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_sharded_checkpoint_roundtrip_and_reshard_world2():
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank, args=(r, port, results)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=180)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}
