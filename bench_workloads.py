"""Workload generators and the on-hardware value check used by bench.py at N > 1.

BASELINE.json configs 2-5 as synthetic shapes: tables, sharding plan, per-rank batch,
length law, id law.  `parity_check` runs ONE seeded training step of the same plan kind (rows capped,
small batch) through ShardedEmbeddingBagCollection on the GPUs and compares it with the CPU oracle
(the oracle is used here as the checker only, never on the timed path):

  routing   the ids / lengths every owner received, bit for bit, against the layout the contract
            states ([feature][source rank][batch]; row-wise: ids of my row block, rebased, in
            original order);
  forward   every rank's pooled rows against oracle.embedding_forward on its local batch;
  update    reassembled tables (one of every sharding kind, first and last) after the fused backward on
            the gathered global batch, against a float64 restatement of the update; the C oracle's own
            fp32 result is measured against the same yardstick and reported next to it.
"""
from __future__ import annotations

import numpy as np

CRITEO = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27, 14992, 5461306, 10,
          5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]


def spec(cfg: int, world: int) -> dict:
    """The full-size workload of BASELINE config `cfg` on `world` GPUs."""
    if cfg == 2:
        return dict(cfg=2, rows=[1_000_000] * 26, D=128, bf16=False, B=16384, plan="balanced", lengths="const", P=20,
                    ids="uniform",
                    name="config 2: 26 tables x 1M rows x dim 128 fp32, pooling factor 20, batch 16384 per GPU, "
                         "uniform ids, SUM pooling, row-wise Adagrad; tables sharded over the GPUs")
    if cfg == 3:
        return dict(cfg=3, rows=list(CRITEO), D=128, bf16=False, B=8192, plan="tw_greedy", lengths="const", P=1,
                    ids="uniform",
                    name="config 3: 26 tables with the Criteo-Kaggle cardinalities (33.8M rows) x dim 128 fp32, "
                         "table-wise on a greedy size-balanced plan, one id per bag, batch 8192 per GPU")
    if cfg == 4:
        return dict(cfg=4, rows=[50_000_000] * 8, D=128, bf16=True, B=8192, plan="rw", lengths="const", P=20,
                    ids="uniform",
                    name="config 4: 8 tables x 50M rows x dim 128 bf16 (102 GB) row-wise over the GPUs, pooling "
                         "factor 20, batch 8192 per GPU, fp32 accumulate and row-wise Adagrad state")
    if cfg == 5:
        return dict(cfg=5, rows=[1_000_000] * 8 + [20_000_000] * 8, D=128, bf16=False, B=65536 // world, plan="mixed8",
                    lengths="zipf500", P=None, ids="zipf",
                    name="config 5: 8 tables x 1M rows table-wise + 8 tables x 20M rows row-wise, dim 128 fp32, "
                         f"global batch 65536 ({65536 // world} per GPU), bag lengths min(zipf(1.5), 500), ids Zipf(1.05)")
    raise ValueError(f"unknown config {cfg}")


def make_tables(t, sp: dict, rows=None):
    rows = sp["rows"] if rows is None else rows
    dt = t.DataType.BF16 if sp["bf16"] else t.DataType.FP32
    return [t.EmbeddingBagConfig(num_embeddings=r, embedding_dim=sp["D"], name=f"t{i}", feature_names=[f"f{i}"],
                                 pooling=t.PoolingType.SUM, data_type=dt) for i, r in enumerate(rows)]


def make_plan(td, sp: dict, tables, world: int, override: str | None = None):
    kind = override or sp["plan"]
    F = len(tables)
    if kind == "balanced":
        return td.balanced_plan(tables, world)
    if kind == "tw":
        return {tb.name: td.table_wise(i % world) for i, tb in enumerate(tables)}
    if kind == "tw_greedy":
        return td.balanced_table_wise_plan(tables, world)
    if kind == "rw":
        return {tb.name: td.row_wise() for tb in tables}
    if kind == "mixed":          # config 2's remainder row-wise
        n_tw = (F // world) * world
        return {tb.name: (td.table_wise(i % world) if i < n_tw else td.row_wise()) for i, tb in enumerate(tables)}
    if kind == "mixed8":         # first half table-wise (round-robin), second half row-wise
        return {tb.name: (td.table_wise(i % world) if i < F // 2 else td.row_wise()) for i, tb in enumerate(tables)}
    raise ValueError(kind)


def np_lengths(sp: dict, rng, n: int, cap: int = 500) -> np.ndarray:
    if sp["lengths"] == "const":
        return np.full(n, sp["P"], np.int32)
    return np.minimum(rng.zipf(1.5, size=n), cap).astype(np.int32)


def np_ids(sp: dict, rng, rows: int, n: int) -> np.ndarray:
    if sp["ids"] == "uniform":
        return rng.integers(0, rows, size=n).astype(np.int64)
    return np.minimum(rng.zipf(1.05, size=n) - 1, rows - 1).astype(np.int64)


def device_batch(torch, sp: dict, rows, B: int, dev, seed: int):
    """One full-size batch on the device: (lengths int32[F*B], ids int64[L])."""
    rng = np.random.default_rng(seed)
    F = len(rows)
    lengths_np = np_lengths(sp, rng, F * B)
    lengths = torch.from_numpy(lengths_np).to(dev)
    per_f = lengths_np.reshape(F, B).sum(axis=1, dtype=np.int64)
    L = int(per_f.sum())
    if sp["ids"] == "uniform":
        gen = torch.Generator(device=dev).manual_seed(seed)
        rows_pos = torch.tensor(rows, dtype=torch.int64, device=dev).repeat_interleave(torch.from_numpy(per_f).to(dev))
        ids = torch.randint(0, 1 << 62, (L,), device=dev, generator=gen, dtype=torch.int64) % rows_pos
    else:
        z = rng.zipf(1.05, size=L) - 1
        rows_pos = np.repeat(np.asarray(rows, dtype=np.int64), per_f)
        ids = torch.from_numpy(np.minimum(z, rows_pos - 1).astype(np.int64)).to(dev)
    return lengths, ids


# --------------------------------------------------------------------------- value check on the GPUs
PARITY_ROW_CAP, PARITY_B, PARITY_LEN_CAP, PARITY_LR, PARITY_SEED = 10_000, 256, 64, 0.05, 11


def _parity_batch(sp, rows, rank):
    rng = np.random.default_rng(7000 + 13 * rank)
    F, B = len(rows), PARITY_B
    lengths = np_lengths(sp, rng, F * B, cap=PARITY_LEN_CAP)
    lengths[rank::97] = 0                                    # some empty bags on every rank
    off = np.concatenate([[0], np.cumsum(lengths, dtype=np.int64)])
    idx = np.concatenate([np_ids(sp, rng, rows[f], int(off[(f + 1) * B] - off[f * B])) for f in range(F)])
    grad = rng.normal(size=(B, F * sp["D"])).astype(np.float32)
    return lengths, idx.astype(np.int64), grad, off


def _expected_landing(m, per, rows, world, rank, B):
    """What rank `rank` must have received: lengths and ids in [local feature][source rank][batch] order."""
    slots = [(f, None) for f in m._tw_feats[rank]] + [(f, None) for (f, _) in m._cw_slots[rank]] + \
            [(f, bs) for f, bs in zip(m._rw_feats, m._rw_block)]
    lens, ids = [], []
    for f, bs in slots:
        for src in range(world):
            ln, idx, _, off = per[src]
            s, e = int(off[f * B]), int(off[(f + 1) * B])
            seg_ids, seg_len = idx[s:e], ln[f * B:(f + 1) * B]
            if bs is None:
                lens.append(seg_len)
                ids.append(seg_ids)
            else:                                            # my row block only, rebased, original order kept
                mine = (seg_ids // bs) == rank
                bag_of = np.repeat(np.arange(B), seg_len)
                lens.append(np.bincount(bag_of[mine], minlength=B).astype(np.int32))
                ids.append(seg_ids[mine] - rank * bs)
    cat = lambda parts, dt: np.concatenate(parts).astype(dt) if parts else np.zeros(0, dt)
    return cat(lens, np.int32), cat(ids, np.int64)


def parity_check(cfg: int, world: int, rank: int, dev, plan_override: str | None = None, kernel_ops=None,
                 peer_memory=None) -> dict:
    """COLLECTIVE.  See the module docstring.  Returns the `parity_check` block of the bench line.
    (`kernel_ops` / `peer_memory`: only tests/ pass them, to run this very check on CPU ranks.)"""
    import torch
    import torch.distributed as dist
    import oracle as orc
    import tdfo_b200 as t
    from tdfo_b200 import distributed as td

    sp = spec(cfg, world)
    D, bf16 = sp["D"], sp["bf16"]
    rows = [min(r, PARITY_ROW_CAP) for r in sp["rows"]]
    F, B = len(rows), PARITY_B
    tables = make_tables(t, sp, rows)
    plan = make_plan(td, sp, tables, world, plan_override)
    m = td.ShardedEmbeddingBagCollection(tables, plan, dev, init_seed=PARITY_SEED, kernel_ops=kernel_ops,
                                         peer_memory=peer_memory)
    sync = torch.cuda.synchronize if dev.type == "cuda" else (lambda: None)
    m.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=PARITY_LR, eps=1e-8)
    m.train()
    per = [_parity_batch(sp, rows, r) for r in range(world)]
    lengths, idx, grad, off = per[rank]
    keys = [f"f{i}" for i in range(F)]
    kjt = t.KeyedJaggedTensor.from_lengths_sync(keys, torch.from_numpy(idx), torch.from_numpy(lengths)).to(dev)
    out = m(kjt).values()
    sync()
    got_out = out.detach().cpu().numpy().copy()

    # ---- routing: what landed on this rank, bit for bit
    routing_ok = None
    if m._use_peer and m._peer is not None:
        z, Fl = m._peer.zones[0], m._F_local
        n_bags = Fl * world * B
        got_len = z.len_x[:n_bags].cpu().numpy()
        total = int(z.off_x[n_bags].item()) if n_bags else 0
        got_ids = z.ids_x[:total].cpu().numpy()
        want_len, want_ids = _expected_landing(m, per, rows, world, rank, B)
        routing_ok = bool(np.array_equal(got_len, want_len) and np.array_equal(got_ids, want_ids))

    out.backward(torch.from_numpy(grad).to(dev))
    sync()
    if dev.type == "cuda":
        m.check_errors()
    full = m.full_state_dict()                               # collective, reassembled on the host

    # ---- oracle on the same data
    init = [td.ShardedEmbeddingBagCollection.initial_table(tb, PARITY_SEED).numpy() for tb in tables]
    w0 = np.concatenate(init)
    w_ref = orc.bf16_bits(w0) if bf16 else w0
    w_old32 = orc.bf16_to_f32(w_ref) if bf16 else w0
    base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
    col = (np.arange(F) * D).astype(np.int64)
    want, _ = orc.embedding_forward(w_ref, D, idx, off, None, base, rows, col, F, B, orc.POOL_SUM)
    mag, _ = orc.embedding_forward(np.abs(w_old32), D, idx, off, None, base, rows, col, F, B, orc.POOL_SUM)
    # relative to the sum of magnitudes pooled into each value (forward-error form of a relative bound)
    err_fwd = float((np.abs(got_out - want) / np.maximum(mag, 1e-30))[mag > 0].max()) if (mag > 0).any() else 0.0

    g_len = np.concatenate([per[r][0].reshape(F, B)[f] for f in range(F) for r in range(world)]).astype(np.int32)
    g_idx = np.concatenate([per[r][1][per[r][3][f * B]:per[r][3][(f + 1) * B]] for f in range(F) for r in range(world)])
    g_grad = np.concatenate([per[r][2] for r in range(world)], axis=0)
    kinds = [plan[tb.name].sharding_type.value for tb in tables]
    check = sorted({0, F - 1} | {kinds.index(k) for k in set(kinds)})[:4]
    g_off = orc.complete_cumsum(g_len)
    GB = world * B
    err_rows = err_oracle = 0.0
    for i in check:
        # Yardstick: a float64 restatement of the update of this table (oracle.fp64_backward_update); the GPU result
        # and the C oracle's fp32 result are both measured against it, relative to |w| + step * sum |gradient terms|
        # (hot rows are summed in another order on the GPU than in the oracle: neither fp32 order is "the" answer).
        r_ = rows[i]
        ps = plan[f"t{i}"]
        k = len(ps.ranks) if ps.sharding_type == td.ShardingType.COLUMN_WISE else 1   # a column shard keeps its own state
        dc = D // k
        ids_i = g_idx[g_off[i * GB]:g_off[(i + 1) * GB]]
        off_i = g_off[i * GB:(i + 1) * GB + 1] - g_off[i * GB]
        old = w_old32[base[i]:base[i] + r_]
        got = full[f"embedding_bags.t{i}.weight"].float().numpy()
        for j in range(k):
            c0 = j * dc
            g_sh = np.ascontiguousarray(g_grad[:, i * D + c0:i * D + c0 + dc])
            old_sh = np.ascontiguousarray(old[:, c0:c0 + dc])
            w64, _, _, _, scale = orc.fp64_backward_update(old_sh, dc, np.zeros(r_), None, g_sh, ids_i, off_i, [0], [r_], [0],
                                                           1, GB, orc.POOL_SUM, orc.OPT_ROWWISE_ADAGRAD, PARITY_LR, 1e-8)
            err_rows = max(err_rows, orc.within_budget(got[:, c0:c0 + dc], w64, scale, 1.0))
            w_o = orc.bf16_bits(old_sh) if bf16 else old_sh.copy()
            orc.embedding_backward_update(w_o, dc, np.zeros(r_, np.float32), None, g_sh, ids_i, off_i, [0], [r_], [0], 1, GB,
                                          orc.POOL_SUM, orc.OptArgs(orc.OPT_ROWWISE_ADAGRAD, PARITY_LR, 1e-8))
            err_oracle = max(err_oracle, orc.within_budget(orc.bf16_to_f32(w_o) if bf16 else w_o, w64, scale, 1.0))
    tol = 1e-2 if bf16 else 1e-5
    stats = torch.tensor([err_fwd, err_rows, 0.0 if routing_ok in (True, None) else 1.0], dtype=torch.float64, device=dev)
    dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    err_fwd, err_rows, bad_routing = (float(x) for x in stats.tolist())
    ok = err_fwd <= tol and err_rows <= tol and bad_routing == 0.0
    del m
    if dev.type == "cuda":
        torch.cuda.empty_cache()
    return {"ok": bool(ok), "max_rel_err_fwd": err_fwd, "max_rel_err_rows": err_rows,
            "oracle_vs_float64_rows": err_oracle,
            "routing_bit_exact": None if routing_ok is None else bad_routing == 0.0, "tolerance": tol,
            "tables_checked": [f"t{i} ({kinds[i]})" for i in check],
            "shape": f"same plan kind as the timed run, rows capped at {PARITY_ROW_CAP}, batch {B} per GPU, one fused "
                     "row-wise Adagrad step; errors are relative to the sum of magnitudes entering each value"}
