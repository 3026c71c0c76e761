"""Scaled-down BASELINE configs 3, 4 and 5 as world_size-2 gloo parity cases.
Same table mix, sharding plan kind, pooling factor / length law, id law, dtype and optimizer as the
config names, with rows, batch and D shrunk so the oracle finishes in seconds:

  3  26 tables with the Criteo-Kaggle cardinalities (/2000), table-wise on a greedy size-balanced
     plan, one id per bag (P = 1), fp32
  4  8 equal tables row-wise, bf16 weights (fp32 accumulate and state), 20 ids per bag
  5  8 small tables table-wise + 8 large ones row-wise, bag lengths min(zipf(1.5), cap) with a few
     bags at the cap, ids Zipf(1.05): long backward segments and unbalanced ranks

Checked per rank for two training steps with row-wise Adagrad: pooled output == unsharded oracle on
the local batch, and the reassembled tables == the unsharded oracle update on the global batch
(1e-5 fp32, 1e-2 bf16).  Routing is host + oracle here; tests/test_gpu_sharded.py runs the same
shapes on the CUDA kernels."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle as orc
from test_sharded_gloo import free_port

W, STEPS, LR = 2, 2, 0.05
CRITEO = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27, 14992, 5461306, 10,
          5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]


def config(cfg, D=16, B=8):
    """-> dict(rows, kinds ('tw' | 'rw' per table), B, D, bf16, lengths(rng, n), ids(rng, rows, n))"""
    c = _config(cfg)
    c.update(D=D, B=B)
    return c


def _config(cfg):
    if cfg == 3:
        rows = [max(3, c // 2000) for c in CRITEO]
        return dict(rows=rows, kinds=None, bf16=False, lengths=lambda rng, n: np.ones(n, np.int32),
                    ids=lambda rng, r, n: rng.integers(0, r, size=n))
    if cfg == 4:
        return dict(rows=[5000] * 8, kinds=["rw"] * 8, bf16=True,
                    lengths=lambda rng, n: np.full(n, 20, np.int32), ids=lambda rng, r, n: rng.integers(0, r, size=n))
    if cfg == 5:
        def lengths(rng, n):
            ln = np.minimum(rng.zipf(1.5, size=n), 60).astype(np.int32)
            ln[rng.integers(0, n, size=max(1, n // 30))] = 60           # a few bags at the cap
            return ln
        return dict(rows=[100] * 8 + [2000] * 8, kinds=["tw"] * 8 + ["rw"] * 8, bf16=False, lengths=lengths,
                    ids=lambda rng, r, n: np.minimum(rng.zipf(1.05, size=n) - 1, r - 1))
    raise ValueError(cfg)


def batch(c, rank, step):
    rng = np.random.default_rng(10_000 * step + 100 * rank + 7)
    F, B = len(c["rows"]), c["B"]
    lengths = c["lengths"](rng, F * B)
    off = orc.complete_cumsum(lengths)
    idx = np.concatenate([c["ids"](rng, c["rows"][f], int(off[(f + 1) * B] - off[f * B])) for f in range(F)])
    grad = rng.normal(size=(B, F * c["D"])).astype(np.float32)
    return lengths, idx.astype(np.int64), grad


def global_batch(c, step):
    F, B = len(c["rows"]), c["B"]
    per = [batch(c, r, step) for r in range(W)]
    offs = [orc.complete_cumsum(p[0]) for p in per]
    lengths = np.concatenate([per[r][0].reshape(F, B)[f] for f in range(F) for r in range(W)]).astype(np.int32)
    idx = np.concatenate([per[r][1][offs[r][f * B]:offs[r][(f + 1) * B]] for f in range(F) for r in range(W)])
    return lengths, idx, np.concatenate([p[2] for p in per], axis=0)


def run_rank(rank, port, cfg, results, cuda=False, D=16, B=8):
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
        c = config(cfg, D, B)
        rows, F = c["rows"], len(c["rows"])
        dt = t.DataType.BF16 if c["bf16"] else t.DataType.FP32
        tables = [t.EmbeddingBagConfig(num_embeddings=r, embedding_dim=D, name=f"t{i}", feature_names=[f"f{i}"],
                                       pooling=t.PoolingType.SUM, data_type=dt) for i, r in enumerate(rows)]
        if c["kinds"] is None:
            plan = td.balanced_table_wise_plan(tables, W)
            load = [sum(r for r, tb in zip(rows, tables) if plan[tb.name].rank == d) for d in range(W)]
            assert max(load) - min(load) <= max(rows)                # greedy balance: within one table
        else:
            tw = [i for i, k in enumerate(c["kinds"]) if k == "tw"]
            plan = {f"t{i}": (td.table_wise(tw.index(i) % W) if k == "tw" else td.row_wise())
                    for i, k in enumerate(c["kinds"])}
        m = td.ShardedEmbeddingBagCollection(tables, plan, dev, kernel_ops=None if cuda else OracleKernelOps(),
                                             init_seed=9)
        m.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=LR, eps=1e-8)
        m.train()

        init = [td.ShardedEmbeddingBagCollection.initial_table(tb, 9).numpy() for tb in tables]
        w_ref = np.concatenate(init)
        if c["bf16"]:
            w_ref = orc.bf16_bits(w_ref)
        h_ref = np.zeros(sum(rows), np.float32)
        base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
        col = (np.arange(F) * D).astype(np.int64)
        rtol, atol = (1e-2, 1e-2) if c["bf16"] else (1e-5, 4e-6 if cuda else 2e-6)
        # Config 5 on the GPU: hot rows are reduced as per-group partial sums (another fp32 association than
        # the oracle's left-to-right sum), so updated rows differ by ~1e-6 relative; a bag that holds the
        # same hot row up to 60 times adds those differences up.  Bound: 1e-5 * bag length * max|w|.
        out_atol = 1e-4 if (cuda and cfg == 5) else atol
        keys = [f"f{i}" for i in range(F)]
        for step in range(STEPS):
            lengths, idx, grad = batch(c, rank, step)
            kjt = t.KeyedJaggedTensor.from_lengths_sync(keys, torch.from_numpy(idx), torch.from_numpy(lengths)).to(dev)
            out = m(kjt).values()
            want, oob = orc.embedding_forward(w_ref, D, idx, orc.complete_cumsum(lengths), None, base, rows, col, F, B,
                                              orc.POOL_SUM)
            assert not oob
            np.testing.assert_allclose(out.detach().cpu().numpy(), want, rtol=rtol, atol=out_atol, err_msg=f"step {step}")
            out.backward(torch.from_numpy(grad).to(dev))
            g_len, g_idx, g_grad = global_batch(c, step)
            orc.embedding_backward_update(w_ref, D, h_ref, None, g_grad, g_idx, orc.complete_cumsum(g_len), base, rows,
                                          col, F, W * B, orc.POOL_SUM, orc.OptArgs(orc.OPT_ROWWISE_ADAGRAD, LR, 1e-8))
        full = m.full_state_dict()
        ref32 = orc.bf16_to_f32(w_ref) if c["bf16"] else w_ref
        for i, r in enumerate(rows):
            got = full[f"embedding_bags.t{i}.weight"].float().numpy()
            np.testing.assert_allclose(got, ref32[base[i]:base[i] + r], rtol=rtol, atol=atol, err_msg=f"table t{i}")
        assert int(m._err_flag.item()) == 0
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("cfg", [3, 4, 5])
def test_scaled_baseline_config_world2(cfg):
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank, args=(r, port, cfg, results)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=240)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}
