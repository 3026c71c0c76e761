"""The sharded path on real GPUs (needs >= 2): same invariant as tests/test_sharded_gloo.py, but
with the CUDA kernels and NCCL.  Skipped on a single-GPU box."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle as orc
from test_sharded_gloo import (B, D, W, free_port, global_batch, local_batch, make_tables, reference_update,
                               run_rank_edges, run_rank_sequence)

pytestmark = pytest.mark.gpu


def run_rank(rank, port, plan_kind, peer, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=W, device_id=dev)
    try:
        import tdfo_b200 as t
        from tdfo_b200 import distributed as td
        tables = make_tables(t)
        rows = [c.num_embeddings for c in tables]
        plan = {"tw": {"t0": td.table_wise(0), "t1": td.table_wise(1), "t2": td.table_wise(0), "t3": td.table_wise(1),
                       "t4": td.table_wise(1)},
                "rw": {c.name: td.row_wise() for c in tables},
                "mixed": {"t0": td.table_wise(1), "t1": td.row_wise(), "t2": td.table_wise(0), "t3": td.row_wise(),
                          "t4": td.table_wise(1)},
                "cw": {"t0": td.column_wise([0, 1]), "t1": td.column_wise([1, 0]), "t2": td.table_wise(0),
                       "t3": td.table_wise(1), "t4": td.column_wise([0, 1])},
                "all3": {"t0": td.column_wise([1, 0]), "t1": td.row_wise(), "t2": td.table_wise(0), "t3": td.row_wise(),
                         "t4": td.column_wise([0, 1])}}[plan_kind]
        m = td.ShardedEmbeddingBagCollection(tables, plan, dev, init_seed=3, peer_memory=peer)
        m.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=0.05, eps=1e-8)
        m.train()
        lengths, idx, grad = local_batch(rank, rows)
        kjt = t.KeyedJaggedTensor.from_lengths_sync([f"f{i}" for i in range(len(rows))], torch.from_numpy(idx),
                                                    torch.from_numpy(lengths)).to(dev)
        out = m(kjt).values()
        out_host = out.detach().cpu().numpy().copy()
        (out * torch.from_numpy(grad).to(dev)).sum().backward()
        torch.cuda.synchronize()
        assert int(m._err_flag.item()) == 0

        full = np.concatenate([td.ShardedEmbeddingBagCollection.initial_table(c, 3).numpy() for c in tables])
        base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
        col = (np.arange(len(rows)) * D).astype(np.int64)
        want, _ = orc.embedding_forward(full, D, idx, orc.complete_cumsum(lengths), None, base, rows, col,
                                        len(rows), B, orc.POOL_SUM)
        np.testing.assert_allclose(out_host, want, rtol=1e-5, atol=1e-6)
        g_len, g_idx, g_grad = global_batch(rows)
        w_ref = reference_update(full, rows, plan, td, g_len, g_idx, g_grad, orc.OPT_ROWWISE_ADAGRAD)
        full_sd = m.full_state_dict()                       # shards reassembled (rows / columns / whole tables)
        for i, c in enumerate(tables):
            got = full_sd[f"embedding_bags.{c.name}.weight"].numpy()
            np.testing.assert_allclose(got, w_ref[base[i]:base[i] + rows[i]], rtol=1e-5, atol=2e-6,
                                       err_msg=f"table {c.name} ({plan_kind})")
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("plan_kind,peer", [("tw", False), ("tw", True), ("rw", False), ("rw", True), ("mixed", False),
                                            ("mixed", True), ("cw", False), ("cw", True), ("all3", True)])
def test_sharded_equals_unsharded_on_2_gpus(plan_kind, peer):
    """peer=True: table-wise pooled rows and gradient rows travel through NVLink peer memory written
    by the kernels themselves; peer=False: NCCL all-to-alls."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank, args=(r, port, plan_kind, peer, results)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=300)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}


@pytest.mark.parametrize("plan_kind", ["tw", "rw", "mixed"])
def test_sharded_sequence_on_2_gpus(plan_kind):
    """ShardedEmbeddingCollection (one row per id) with the CUDA kernels over NCCL."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank_sequence, args=(r, port, plan_kind, results, True)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=300)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}


@pytest.mark.parametrize("cfg", [3, 4, 5])
def test_scaled_baseline_configs_on_2_gpus(cfg):
    """BASELINE configs 3 (Criteo cardinalities, table-wise, P=1), 4 (row-wise bf16, P=20) and 5 (mixed plan,
    Zipf lengths and ids) scaled down, two training steps on the CUDA kernels with D = 128
    (tests/test_configs_gloo.py holds the checks)."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from test_configs_gloo import run_rank as run_config_rank
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_config_rank, args=(r, port, cfg, results, True, 128, 48)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=300)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}


def run_rank_replicated_adapter(rank, port, results):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=W, device_id=dev)
    try:
        import tdfo_b200 as t
        from tdfo_b200.adapters import DLPackEmbeddingTables
        tables = {"user_id": (90, 16), "item_id": (120, 16)}
        Bn, lr = 32, 0.1
        tabs = DLPackEmbeddingTables(tables, dev, optimizer=t.EmbOptimType.EXACT_SGD, learning_rate=lr)
        w = {n: np.random.default_rng(3).uniform(-1, 1, size=s).astype(np.float32) for n, s in tables.items()}
        tabs.load_state_dict({f"embedding_bags.{n}.weight": torch.from_numpy(a) for n, a in w.items()})   # same replica everywhere
        per = []
        for r in range(W):
            rng = np.random.default_rng(50 + r)
            per.append(({n: rng.integers(0, s[0], size=Bn) for n, s in tables.items()},
                        {n: rng.normal(size=(Bn, s[1])).astype(np.float32) for n, s in tables.items()}))
        ids, grads = per[rank]
        tabs.apply_gradients({n: torch.from_numpy(a).to(dev) for n, a in ids.items()},
                             {n: torch.from_numpy(g).to(dev) for n, g in grads.items()}, replicated=True, average=True)
        for n, (rows, d) in tables.items():       # pmean + SGD restated: global batch in rank order, gradients / W
            g_ids = np.concatenate([per[r][0][n] for r in range(W)]).astype(np.int64)
            g_grad = np.concatenate([per[r][1][n] for r in range(W)]) / W
            orc.embedding_backward_update(w[n], d, None, None, g_grad.astype(np.float32), g_ids,
                                          np.arange(W * Bn + 1, dtype=np.int64), [0], [rows], [0], 1, W * Bn,
                                          orc.POOL_SUM, orc.OptArgs(orc.OPT_SGD, lr))
            got = tabs.state_dict()[f"embedding_bags.{n}.weight"].cpu().numpy()
            np.testing.assert_allclose(got, w[n], rtol=1e-5, atol=2e-6)
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_replicated_adapter_tables_stay_identical_on_2_gpus():
    """jax pmap / tf MirroredStrategy shape: every rank holds a replica and feeds its own batch; the sparse
    apply_gradients gathers (ids, rows) so all replicas take the same step (jax-flax/train_dp.py:63-65)."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank_replicated_adapter, args=(r, port, results)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=300)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}


def test_empty_and_lopsided_batches_on_2_gpus():
    """The peer-memory exchange with ranks that send nothing, steps without any id, one live feature."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank_edges, args=(r, port, results, True)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=200)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}


@pytest.mark.parametrize("case,prefetch", [("w4", False), ("w4", True), ("w4_all_rw", False), ("w8_bench_plan", False),
                                           ("w8_mixed", False), ("w8_mixed", True)])
def test_peer_memory_exchange_on_4_and_8_gpus(case, prefetch):
    """The cases of tests/test_peer_layout_gloo.py (uneven table-wise ownership + column shards + a row-wise
    table on 4 ranks; bench.py's balanced plan, BASELINE config 4's all-row-wise and config 5's mixed plan on
    4 / 8 ranks) on real symmetric memory over NVLink: two training steps (four with the input dist
    prefetched), every rank's pooled rows against the oracle on its batch, all tables afterwards against the
    oracle's update on the global batch."""
    from test_peer_layout_gloo import CASES, run_rank as run_layout_rank
    world = CASES[case]["W"]
    if not torch.cuda.is_available() or torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_layout_rank, args=(r, port, "", case, results, prefetch, True)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=300)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {r: "ok" for r in range(world)}


def test_dmp_over_the_reference_item_table_on_2_gpus():
    """torchrec/train.py:235-254 on two GPUs: DistributedModelParallel shards the item EmbeddingCollection row-wise with
    fused ADAM (the reference's GPU choice), the model calls jagged_2d_to_dense / to_padded_dense as
    torchrec/models.py:163-178,208-219 does, the dense half is under DDP; one step equals the unsharded oracle."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from test_dmp_gloo import run_rank_sequence as run_dmp_rank
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_dmp_rank, args=(r, port, results, True)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=300)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}
