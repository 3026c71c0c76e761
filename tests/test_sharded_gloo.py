"""World-size-2 (gloo, CPU) test of the sharded path's host logic: input dist (permute /
bucketise / two all-to-alls / recv permute), output dist (all-to-all for table-wise, ordered
partial-sum exchange for row-wise) and the reverse exchanges in backward.

Invariant (the one torchrec's own sharding tests use): every rank's sharded
output equals the unsharded lookup of its local batch, and after one backward the gathered
shards equal the unsharded fused update applied to the global batch.  The compute kernels are
replaced by the oracle-backed table from tests/oracle_ops.py, so this covers routing only;
the same test runs the CUDA kernels on 2 GPUs in tests/test_gpu_sharded.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle as orc

W, B, D = 2, 6, 8


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def make_tables(t):
    cfg = lambda name, rows, feats: t.EmbeddingBagConfig(num_embeddings=rows, embedding_dim=D, name=name,
                                                         feature_names=feats, pooling=t.PoolingType.SUM)
    return [cfg("t0", 50, ["f0"]), cfg("t1", 41, ["f1"]), cfg("t2", 30, ["f2"]), cfg("t3", 17, ["f3"]),
            cfg("t4", 9, ["f4"])]


def local_batch(rank, rows):
    rng = np.random.default_rng(100 + rank)
    F = len(rows)
    lengths = rng.integers(0, 5, size=F * B).astype(np.int32)
    lengths[rank::7] = 0
    offsets = orc.complete_cumsum(lengths)
    idx = np.concatenate([rng.integers(0, rows[f], size=int(offsets[(f + 1) * B] - offsets[f * B])) for f in range(F)])
    grad = rng.normal(size=(B, F * D)).astype(np.float32)
    return lengths, idx.astype(np.int64), grad


def global_batch(rows):
    """Concatenate the ranks' batches per feature: sample index = rank * B + b."""
    F = len(rows)
    per = [local_batch(r, rows) for r in range(W)]
    lengths = np.concatenate([per[r][0].reshape(F, B)[f] for f in range(F) for r in range(W)]).astype(np.int32)
    idx_parts = []
    for f in range(F):
        for r in range(W):
            off = orc.complete_cumsum(per[r][0])
            idx_parts.append(per[r][1][off[f * B]:off[(f + 1) * B]])
    grad = np.concatenate([per[r][2] for r in range(W)], axis=0)
    return lengths, np.concatenate(idx_parts), grad


def reference_update(full, rows, plan, td, g_len, g_idx, g_grad, code, lr=0.05):
    """Unsharded oracle update on the global batch, table by table.  A column-wise table is updated
    shard by shard: each column shard is an independent table of D/k columns ([UPSTREAM] torchrec
    column-wise shards are separate TBE tables), which matters only for row-wise Adagrad, whose
    per-row state then averages g^2 over the shard's own columns."""
    w_ref = full.copy()
    base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
    g_off = orc.complete_cumsum(g_len)
    GB = W * B
    for i, r_ in enumerate(rows):
        ps = plan[f"t{i}"]
        k = len(ps.ranks) if ps.sharding_type == td.ShardingType.COLUMN_WISE else 1
        dc = D // k
        idx = g_idx[g_off[i * GB]:g_off[(i + 1) * GB]]
        off = g_off[i * GB:(i + 1) * GB + 1] - g_off[i * GB]
        for j in range(k):
            c0 = j * dc
            w = np.ascontiguousarray(w_ref[base[i]:base[i] + r_, c0:c0 + dc])
            g = np.ascontiguousarray(g_grad[:, i * D + c0:i * D + c0 + dc])
            orc.embedding_backward_update(w, dc, np.zeros(r_, np.float32), None, g, idx, off, [0], [r_], [0], 1, GB,
                                          orc.POOL_SUM, orc.OptArgs(code, lr, 1e-8))
            w_ref[base[i]:base[i] + r_, c0:c0 + dc] = w
    return w_ref


def run_rank(rank, port, plan_kind, optimizer_name, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=W)
    try:
        import tdfo_b200 as t
        from tdfo_b200 import distributed as td
        from oracle_ops import OracleKernelOps
        tables = make_tables(t)
        rows = [c.num_embeddings for c in tables]
        if plan_kind == "tw":
            plan = {"t0": td.table_wise(0), "t1": td.table_wise(1), "t2": td.table_wise(0), "t3": td.table_wise(1),
                    "t4": td.table_wise(1)}
        elif plan_kind == "rw":
            plan = {c.name: td.row_wise() for c in tables}
        elif plan_kind == "cw":       # column shards (D/2 columns each) next to whole tables
            plan = {"t0": td.column_wise([0, 1]), "t1": td.column_wise([1, 0]), "t2": td.table_wise(0),
                    "t3": td.table_wise(1), "t4": td.column_wise([0, 1])}
        elif plan_kind == "all3":     # table-wise + column-wise + row-wise in one module
            plan = {"t0": td.column_wise([1, 0]), "t1": td.row_wise(), "t2": td.table_wise(0), "t3": td.row_wise(),
                    "t4": td.column_wise([0, 1])}
        else:
            plan = {"t0": td.table_wise(1), "t1": td.row_wise(), "t2": td.table_wise(0), "t3": td.row_wise(),
                    "t4": td.table_wise(1)}
        m = td.ShardedEmbeddingBagCollection(tables, plan, torch.device("cpu"), kernel_ops=OracleKernelOps(), init_seed=3)
        opt_type = getattr(t.EmbOptimType, optimizer_name)
        m.set_fused_optimizer(opt_type, learning_rate=0.05, eps=1e-8)
        m.train()
        lengths, idx, grad = local_batch(rank, rows)
        kjt = t.KeyedJaggedTensor.from_lengths_sync([f"f{i}" for i in range(len(rows))], torch.from_numpy(idx),
                                                    torch.from_numpy(lengths))
        out = m(kjt).values()
        (out * torch.from_numpy(grad)).sum().backward()

        # ---- unsharded oracle on the same data
        full = np.concatenate([td.ShardedEmbeddingBagCollection.initial_table(c, 3).numpy() for c in tables])
        base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
        col = (np.arange(len(rows)) * D).astype(np.int64)
        want, _ = orc.embedding_forward(full, D, idx, orc.complete_cumsum(lengths), None, base, rows, col,
                                        len(rows), B, orc.POOL_SUM)
        np.testing.assert_allclose(out.detach().numpy(), want, rtol=1e-5, atol=1e-6)

        g_len, g_idx, g_grad = global_batch(rows)
        code = {"EXACT_SGD": orc.OPT_SGD, "EXACT_ROWWISE_ADAGRAD": orc.OPT_ROWWISE_ADAGRAD}[optimizer_name]
        w_ref = reference_update(full, rows, plan, td, g_len, g_idx, g_grad, code)
        full_sd = m.full_state_dict()                       # shards reassembled (rows / columns / whole tables)
        for i, c in enumerate(tables):
            got = full_sd[f"embedding_bags.{c.name}.weight"].numpy()
            np.testing.assert_allclose(got, w_ref[base[i]:base[i] + rows[i]], rtol=1e-5, atol=1e-6,
                                       err_msg=f"table {c.name} ({plan_kind})")
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("plan_kind,optimizer_name", [("tw", "EXACT_SGD"), ("rw", "EXACT_ROWWISE_ADAGRAD"),
                                                      ("mixed", "EXACT_ROWWISE_ADAGRAD"),
                                                      ("cw", "EXACT_ROWWISE_ADAGRAD"), ("all3", "EXACT_SGD")])
def test_sharded_equals_unsharded_world2(plan_kind, optimizer_name):
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank, args=(r, port, plan_kind, optimizer_name, results)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=180)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}


# ----------------------------------------------------------------------------- sequence (EmbeddingCollection)
def seq_grad(rank, L):
    return np.random.default_rng(500 + rank).normal(size=(L, D)).astype(np.float32)


def run_rank_sequence(rank, port, plan_kind, results, cuda=False, fake_tag=None):
    """cuda: real GPUs (peer memory by default).  fake_tag: CPU ranks driving the peer-memory branch against
    tests/fake_peer.py.  Neither: the all-to-all branch with the oracle-backed kernel table."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    if cuda:
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=W, device_id=dev)
    else:
        dev = torch.device("cpu")
        dist.init_process_group("gloo", rank=rank, world_size=W)
    try:
        import tdfo_b200 as t
        from tdfo_b200 import distributed as td
        from oracle_ops import OracleKernelOps
        specs = [("t0", 50, ["f0"]), ("t1", 41, ["f1"]), ("t2", 30, ["f2"]), ("t3", 17, ["f3"]), ("t4", 9, ["f4"])]
        tables = [t.EmbeddingConfig(num_embeddings=r_, embedding_dim=D, name=n, feature_names=f) for n, r_, f in specs]
        rows = [c.num_embeddings for c in tables]
        F = len(rows)
        plan = {"tw": {"t0": td.table_wise(0), "t1": td.table_wise(1), "t2": td.table_wise(0), "t3": td.table_wise(1),
                       "t4": td.table_wise(1)},
                "rw": {c.name: td.row_wise() for c in tables},
                "mixed": {"t0": td.table_wise(1), "t1": td.row_wise(), "t2": td.table_wise(0), "t3": td.row_wise(),
                          "t4": td.table_wise(1)}}[plan_kind]
        mem = None
        if fake_tag:
            from fake_peer import FakeMem, FakePeerOps
            mem = FakeMem(fake_tag, rank, W)
            m = td.ShardedEmbeddingCollection(tables, plan, dev, kernel_ops=FakePeerOps(mem), init_seed=5, peer_memory=True)
        else:
            m = td.ShardedEmbeddingCollection(tables, plan, dev, kernel_ops=None if cuda else OracleKernelOps(),
                                              init_seed=5)
        assert m._use_peer == bool(cuda or fake_tag)
        m.set_fused_optimizer(t.EmbOptimType.EXACT_SGD, learning_rate=0.05)
        m.train()
        lengths, idx, _ = local_batch(rank, rows)
        kjt = t.KeyedJaggedTensor.from_lengths_sync([f"f{i}" for i in range(F)], torch.from_numpy(idx),
                                                    torch.from_numpy(lengths)).to(dev)
        jts = m(kjt)
        got = torch.cat([jts[f"f{i}"].values() for i in range(F)])
        grad = seq_grad(rank, idx.size)
        (got * torch.from_numpy(grad).to(dev)).sum().backward()

        full = np.concatenate([td.ShardedEmbeddingBagCollection.initial_table(c, 5).numpy() for c in tables])
        base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
        off = orc.complete_cumsum(lengths)
        want, _ = orc.embedding_forward(full, D, idx, off, None, base, rows, None, F, B, orc.POOL_NONE)
        np.testing.assert_array_equal(got.detach().cpu().numpy(), want)      # pure gather + routing: bit-exact
        for i in range(F):                                                  # per-key jagged views
            np.testing.assert_array_equal(jts[f"f{i}"].lengths().cpu().numpy(), lengths[i * B:(i + 1) * B])
            np.testing.assert_array_equal(jts[f"f{i}"].values().detach().cpu().numpy(), want[off[i * B]:off[(i + 1) * B]])

        # unsharded update on the global batch: per feature, rank 0's ids then rank 1's
        per = [local_batch(r_, rows) for r_ in range(W)]
        g_len, g_idx, _ = global_batch(rows)
        g_grad_parts = []
        for f in range(F):
            for r_ in range(W):
                o = orc.complete_cumsum(per[r_][0])
                g_grad_parts.append(seq_grad(r_, per[r_][1].size)[o[f * B]:o[(f + 1) * B]])
        w_ref = full.copy()
        orc.embedding_backward_update(w_ref, D, None, None, np.concatenate(g_grad_parts), g_idx,
                                      orc.complete_cumsum(g_len), base, rows, None, F, W * B, orc.POOL_NONE,
                                      orc.OptArgs(orc.OPT_SGD, 0.05))
        mine = {c.name: m.local_table(c.name).detach().cpu().numpy().copy() for c in tables}
        gathered = [None] * W
        dist.all_gather_object(gathered, mine)
        for i, c in enumerate(tables):
            ps = plan[c.name]
            got_t = gathered[ps.rank][c.name] if ps.sharding_type == td.ShardingType.TABLE_WISE else \
                np.concatenate([gathered[r_][c.name] for r_ in range(W)])
            np.testing.assert_allclose(got_t, w_ref[base[i]:base[i] + rows[i]], rtol=1e-5, atol=1e-6,
                                       err_msg=f"table {c.name} ({plan_kind})")
        results[rank] = "ok"
        if mem is not None:
            dist.barrier()
            mem.cleanup()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("plan_kind,fake_peer", [("tw", False), ("rw", False), ("mixed", False),
                                                 ("tw", True), ("rw", True), ("mixed", True)])
def test_sharded_sequence_equals_unsharded_world2(plan_kind, fake_peer):
    """fake_peer: the peer-memory branch (rows stored by the owners at the ids' own positions, gradient rows pushed
    back) on CPU ranks against tests/fake_peer.py; otherwise the all-to-all branch."""
    import uuid
    port = free_port()
    tag = ("tdfo_fake_" + uuid.uuid4().hex[:8]) if fake_peer else None
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank_sequence, args=(r, port, plan_kind, results, False, tag)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=180)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}


# --------------------------------------------------------------------------- empty and lopsided batches
def run_rank_edges(rank, port, results, cuda=False, prefetch=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    if cuda:
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=W, device_id=dev)
    else:
        dev = torch.device("cpu")
        dist.init_process_group("gloo", rank=rank, world_size=W)
    try:
        import tdfo_b200 as t
        from tdfo_b200 import distributed as td
        from oracle_ops import OracleKernelOps
        tables = make_tables(t)
        rows = [c.num_embeddings for c in tables]
        F = len(rows)
        plan = {"t0": td.column_wise([1, 0]), "t1": td.row_wise(), "t2": td.table_wise(0), "t3": td.row_wise(),
                "t4": td.column_wise([0, 1])}
        m = td.ShardedEmbeddingBagCollection(tables, plan, dev, kernel_ops=None if cuda else OracleKernelOps(),
                                             init_seed=3)
        m.set_fused_optimizer(t.EmbOptimType.EXACT_SGD, learning_rate=0.05)
        m.train()
        rng = np.random.default_rng(rank)
        base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
        col = (np.arange(F) * D).astype(np.int64)
        cases = ("one rank sends nothing", "nobody sends anything", "normal", "one feature only")

        def make(case):
            lengths = rng.integers(0, 4, size=F * B).astype(np.int32)
            if (case == "one rank sends nothing" and rank == 1) or case == "nobody sends anything":
                lengths[:] = 0
            if case == "one feature only":
                lengths[B:] = 0
            off = orc.complete_cumsum(lengths)
            idx = np.concatenate([rng.integers(0, rows[f], size=int(off[(f + 1) * B] - off[f * B]))
                                  for f in range(F)]).astype(np.int64)
            kjt = t.KeyedJaggedTensor.from_lengths_sync([f"f{i}" for i in range(F)], torch.from_numpy(idx),
                                                        torch.from_numpy(lengths)).to(dev)
            return lengths, off, idx, kjt

        nxt = make(cases[0])
        if prefetch:
            m.prefetch_input_dist(nxt[3])
        for ci, case in enumerate(cases):
            lengths, off, idx, kjt = nxt
            full = m.full_state_dict()
            w = np.concatenate([full[f"embedding_bags.t{i}.weight"].numpy() for i in range(F)])
            out = m(kjt).values()
            if ci + 1 < len(cases):
                nxt = make(cases[ci + 1])
                if prefetch:                                     # next batch's exchange before this batch's backward
                    m.prefetch_input_dist(nxt[3])
            want, _ = orc.embedding_forward(w, D, idx, off, None, base, rows, col, F, B, orc.POOL_SUM)
            np.testing.assert_allclose(out.detach().cpu().numpy(), want, rtol=1e-5, atol=1e-6, err_msg=case)
            out.sum().backward()                                 # must not hang or fail on empty exchanges
        if cuda:
            torch.cuda.synchronize()
            assert int(m._err_flag.item()) == 0
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("prefetch", [False, True])
def test_sharded_empty_and_lopsided_batches_world2(prefetch):
    """Ranks with no ids at all, a step where nobody has ids, a single live feature: the exchanges carry
    zero-sized splits and the lookups zero bags (the reference's edge cases are torchrec's).
    prefetch=True issues every input dist through prefetch_input_dist, one step ahead."""
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank_edges, args=(r, port, results, False, prefetch)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=240)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}
