"""BASELINE config 1: embedding-bag + MLP recommender on synthetic MovieLens-1M-shaped data,
CPU / gloo, world_size = 2 (the one configuration of the reference that runs without a GPU:
torchrec/README.md:53, torchrec/train.py:192-198, 276-290).

Each rank trains ShardedEmbeddingBagCollection (mixed table-wise / row-wise plan, fused SGD, the
CPU choice at torchrec/train.py:195) + a DDP-wrapped MLP(64, 32, 1) with BCE for 20 steps on its
own batches of 16 (torchrec/config.toml:14).  The loss curve of every rank must equal, within
1e-5, a single-process restatement that runs the unsharded oracle on the global batch:
dense gradients averaged over ranks (what DDP does), embedding gradients summed over ranks (what
the sharded backward does: each owner applies every rank's gradient rows).

The compute kernels are the oracle-backed table of tests/oracle_ops.py, so this exercises the
routing, the collectives, autograd plumbing and optimizer hand-off — not the CUDA kernels."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

import oracle as orc
from test_sharded_gloo import free_port

W, B, D, STEPS, LR = 2, 16, 16, 20, 0.05
# (table, rows, features, max bag length)
TABLES = [("user", 6040, ["user"], 1), ("movie", 3952, ["movie", "history"], 20), ("gender", 2, ["gender"], 1),
          ("age", 7, ["age"], 1), ("occupation", 21, ["occupation"], 1), ("zip", 3439, ["zip"], 1),
          ("genre", 18, ["genres"], 6)]
FEATURES = [f for _, _, fs, _ in TABLES for f in fs]
FEAT_ROWS = [r for _, r, fs, _ in TABLES for _ in fs]
FEAT_MAXLEN = {"movie": 1, "history": 20, "genres": 6}


def batch(rank, step):
    rng = np.random.default_rng(1000 * step + rank)
    lengths, ids = [], []
    for f, rows in zip(FEATURES, FEAT_ROWS):
        mx = FEAT_MAXLEN.get(f, 1)
        ln = np.ones(B, dtype=np.int32) if mx == 1 else rng.integers(1, mx + 1, size=B).astype(np.int32)
        lengths.append(ln)
        ids.append(rng.integers(0, rows, size=int(ln.sum())))
    labels = rng.integers(0, 2, size=(B, 1)).astype(np.float32)
    return np.concatenate(lengths), np.concatenate(ids).astype(np.int64), labels


def make_mlp():
    torch.manual_seed(7)
    return nn.Sequential(nn.Linear(len(FEATURES) * D, 64), nn.ReLU(), nn.Linear(64, 32), nn.ReLU(), nn.Linear(32, 1))


def reference_losses(tables_full):
    """Single process, unsharded oracle, global batch; returns losses[step][rank]."""
    F = len(FEATURES)
    table_rows = [r for _, r, _, _ in TABLES]
    table_base = np.concatenate([[0], np.cumsum(table_rows)[:-1]])
    feat_base, feat_rows = [], []
    for (_, rows, fs, _), b0 in zip(TABLES, table_base):
        feat_base += [b0] * len(fs)
        feat_rows += [rows] * len(fs)
    col = (np.arange(F) * D).astype(np.int64)
    w = np.concatenate(tables_full).copy()
    mlp = make_mlp()
    opt = torch.optim.SGD(mlp.parameters(), lr=LR)
    bce = nn.BCEWithLogitsLoss()
    curve = []
    for step in range(STEPS):
        per = [batch(r, step) for r in range(W)]
        # global KJT: per feature, rank 0's bags then rank 1's
        g_len = np.concatenate([per[r][0].reshape(F, B)[f] for f in range(F) for r in range(W)]).astype(np.int32)
        parts = []
        for f in range(F):
            for r in range(W):
                off = orc.complete_cumsum(per[r][0])
                parts.append(per[r][1][off[f * B]:off[(f + 1) * B]])
        g_idx = np.concatenate(parts)
        g_off = orc.complete_cumsum(g_len)
        out, _ = orc.embedding_forward(w, D, g_idx, g_off, None, feat_base, feat_rows, col, F, W * B, orc.POOL_SUM)
        x = torch.from_numpy(out).requires_grad_(True)
        logits = mlp(x)
        losses = [bce(logits[r * B:(r + 1) * B], torch.from_numpy(per[r][2])) for r in range(W)]
        opt.zero_grad()
        sum(losses).backward()
        for p in mlp.parameters():
            p.grad /= W                                   # DDP averages dense gradients
        opt.step()
        orc.embedding_backward_update(w, D, None, None, x.grad.numpy(), g_idx, g_off, feat_base, feat_rows, col,
                                      F, W * B, orc.POOL_SUM, orc.OptArgs(orc.OPT_SGD, LR))
        curve.append([float(l) for l in losses])
    return curve


def run_rank(rank, port, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=W)
    try:
        import tdfo_b200 as t
        from tdfo_b200 import distributed as td
        from oracle_ops import OracleKernelOps
        tables = [t.EmbeddingBagConfig(num_embeddings=rows, embedding_dim=D, name=name, feature_names=fs,
                                       pooling=t.PoolingType.SUM) for name, rows, fs, _ in TABLES]
        plan = {"user": td.row_wise(), "movie": td.table_wise(0), "gender": td.table_wise(1), "age": td.table_wise(0),
                "occupation": td.table_wise(1), "zip": td.table_wise(1), "genre": td.row_wise()}
        ebc = td.ShardedEmbeddingBagCollection(tables, plan, torch.device("cpu"), kernel_ops=OracleKernelOps(),
                                               init_seed=11)
        fused = ebc.set_fused_optimizer(t.EmbOptimType.SGD, learning_rate=LR)       # CPU choice, train.py:195
        mlp = nn.parallel.DistributedDataParallel(make_mlp())
        dense = t.KeyedOptimizerWrapper(dict(t.in_backward_optimizer_filter(mlp.named_parameters())),
                                        lambda params: torch.optim.SGD(params, lr=LR))
        optimizer = t.CombinedOptimizer([fused, dense])
        assert not list(t.in_backward_optimizer_filter(ebc.named_parameters()))
        bce = nn.BCEWithLogitsLoss()
        ebc.train()
        mine = []
        for step in range(STEPS):
            lengths, idx, labels = batch(rank, step)
            kjt = t.KeyedJaggedTensor.from_lengths_sync(FEATURES, torch.from_numpy(idx), torch.from_numpy(lengths))
            pooled = ebc(kjt)
            assert pooled.keys() == FEATURES
            loss = bce(mlp(pooled.values()), torch.from_numpy(labels))
            optimizer.zero_grad()
            loss.backward()
            optimizer.step()
            mine.append(loss.item())
        full = [td.ShardedEmbeddingBagCollection.initial_table(c, 11).numpy() for c in tables]
        want = reference_losses(full)
        np.testing.assert_allclose(mine, [want[s][rank] for s in range(STEPS)], rtol=1e-5, atol=1e-6)
        assert abs(mine[0] - mine[-1]) > 1e-4                  # it actually trains
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_config1_loss_curve_world2_equals_unsharded():
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank, args=(r, port, results)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=300)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}
