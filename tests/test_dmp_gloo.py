"""`DistributedModelParallel(module, device, sharders=[...])` (torchrec/train.py:241-247) on two gloo ranks:
the EmbeddingBagCollection inside the model is replaced by its sharded form (table-wise + column-wise +
row-wise plan), which ADOPTS the tables the unsharded module held; dense parameters go under DDP; the
sharder's fused_params reach the fused optimizer; one training step equals the unsharded oracle."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

import oracle as orc
from test_sharded_gloo import B, D, W, free_port, global_batch, local_batch, make_tables, reference_update


class Model(nn.Module):
    def __init__(self, t, tables):
        super().__init__()
        self.sparse = t.EmbeddingBagCollection(tables, device=torch.device("cpu"))
        self.over = nn.Linear(len(tables) * D, 1)

    def forward(self, kjt):
        return self.over(self.sparse(kjt).values())


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
        torch.manual_seed(1234)                               # every rank builds the same model
        model = Model(t, tables)
        held = {k: v.clone() for k, v in model.sparse.state_dict().items()}
        plan = {"t0": td.column_wise([1, 0]), "t1": td.row_wise(), "t2": td.table_wise(0), "t3": td.table_wise(1),
                "t4": td.column_wise([0, 1])}
        fused_params = {"optimizer": t.EmbOptimType.EXACT_SGD, "learning_rate": 0.05}
        dmp = t.DistributedModelParallel(model, device=torch.device("cpu"),
                                         sharders=[t.EmbeddingBagCollectionSharder(fused_params)], plan=plan,
                                         kernel_ops=OracleKernelOps())
        inner = dmp.module.module if isinstance(dmp.module, nn.parallel.DistributedDataParallel) else dmp.module
        assert isinstance(dmp.module, nn.parallel.DistributedDataParallel)          # dense part is data-parallel
        assert isinstance(inner.sparse, td.ShardedEmbeddingBagCollection)
        # the tables came along, split by the plan
        full = inner.sparse.full_state_dict()
        for k, v in held.items():
            torch.testing.assert_close(full[k], v, rtol=0, atol=0)
        # optimizer plumbing of torchrec/train.py:248-254
        dense_named = dict(t.in_backward_optimizer_filter(dmp.named_parameters()))
        assert sorted(n.split(".")[-2] + "." + n.split(".")[-1] for n in dense_named) == ["over.bias", "over.weight"]
        dense_opt = t.KeyedOptimizerWrapper(dense_named, lambda params: torch.optim.SGD(params, lr=0.1))
        optimizer = t.CombinedOptimizer([dmp.fused_optimizer, dense_opt])
        assert dmp.fused_optimizer.param_groups[0]["lr"] == 0.05

        dmp.train()
        lengths, idx, _ = local_batch(rank, rows)
        kjt = t.KeyedJaggedTensor.from_lengths_sync([f"f{i}" for i in range(len(rows))], torch.from_numpy(idx),
                                                    torch.from_numpy(lengths))
        w_over, b_over = inner.over.weight.detach().clone(), inner.over.bias.detach().clone()
        out = dmp(kjt)
        optimizer.zero_grad()
        out.sum().backward()
        optimizer.step()

        # ---- unsharded restatement: pooled rows, then d(sum)/d(pooled) = over.weight for every sample
        w_full = np.concatenate([held[f"embedding_bags.t{i}.weight"].numpy() for i in range(len(rows))])
        base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
        col = (np.arange(len(rows)) * D).astype(np.int64)
        pooled, _ = orc.embedding_forward(w_full, D, idx, orc.complete_cumsum(lengths), None, base, rows, col, len(rows), B,
                                          orc.POOL_SUM)
        np.testing.assert_allclose(out.detach().numpy(), pooled @ w_over.numpy().T + b_over.numpy(), rtol=1e-5, atol=1e-5)
        g_len, g_idx, _ = global_batch(rows)
        g_grad = np.tile(w_over.numpy(), (W * B, 1)).astype(np.float32)
        w_ref = reference_update(w_full, rows, plan, td, g_len, g_idx, g_grad, orc.OPT_SGD, lr=0.05)
        after = inner.sparse.full_state_dict()
        for i, r_ in enumerate(rows):
            np.testing.assert_allclose(after[f"embedding_bags.t{i}.weight"].numpy(), w_ref[base[i]:base[i] + r_],
                                       rtol=1e-5, atol=1e-6, err_msg=f"t{i}")
        assert not torch.equal(inner.over.weight.detach(), w_over)                   # the dense half stepped too
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_dmp_shards_adopts_tables_and_trains_world2():
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank, args=(r, port, results)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=240)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}


# --------------------------------------------------------------------------- the reference's own shape:
# one item table in an EmbeddingCollection, sharded by DMP with fused ADAM (torchrec/train.py:235-254)
V2, T2, B2 = 40, 5, 6


class TinyBert(nn.Module):
    """The reference's HistoryArch + output layer in miniature (torchrec/models.py:150-178, 206-223): item table in
    an EmbeddingCollection, jagged -> [B, T, D] through jagged_2d_to_dense, padding mask from to_padded_dense."""

    def __init__(self, t, device=torch.device("cpu")):
        super().__init__()
        self._t, self._dense_op = t, device.type == "cuda"
        self.embed_collection = t.EmbeddingCollection(
            tables=[t.EmbeddingConfig(name="item_embedding", embedding_dim=D, num_embeddings=V2, feature_names=["item"],
                                      weight_init_max=1.0, weight_init_min=-1.0)], device=device)
        self.out = nn.Linear(D, V2)

    def forward(self, kjt):
        jt = self.embed_collection(kjt)["item"]
        if self._dense_op:                                     # the reference's own calls, on the GPU kernels
            padded = self._t.jagged_2d_to_dense(values=jt.values(), offsets=jt.offsets(), max_sequence_length=T2)
            mask = kjt["item"].to_padded_dense(desired_length=T2, padding_value=0) != 0
            assert bool(mask.all())                            # ids start at 1 in this test: nothing is padding
            return self.out(padded.view(-1, T2, D))
        return self.out(jt.values().view(-1, T2, D))          # every sequence has T2 ids here


def run_rank_sequence(rank, port, results, cuda=False):
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
        torch.manual_seed(77)
        model = TinyBert(t, dev)
        table0 = model.embed_collection.state_dict()["embeddings.item_embedding.weight"].clone().cpu()
        fused_params = {"optimizer": t.EmbOptimType.ADAM, "learning_rate": 3e-2, "weight_decay": 1e-4}
        dmp = t.DistributedModelParallel(module=model, device=dev,
                                         sharders=[t.EmbeddingCollectionSharder(fused_params)],
                                         plan={"item_embedding": td.row_wise()},
                                         kernel_ops=None if cuda else OracleKernelOps())
        inner = dmp.module.module
        assert isinstance(inner.embed_collection, td.ShardedEmbeddingCollection)
        assert inner.embed_collection._use_peer == cuda           # on GPUs: rows and gradients through NVLink peer memory
        dense_optimizer = t.KeyedOptimizerWrapper(dict(t.in_backward_optimizer_filter(dmp.named_parameters())),
                                                  lambda params: torch.optim.Adam(params, lr=3e-2))
        optimizer = t.CombinedOptimizer([dmp.fused_optimizer, dense_optimizer])
        assert sorted(dense_optimizer.params) == ["module.module.out.bias", "module.module.out.weight"]
        sd_key = "module.module.embed_collection.embeddings.item_embedding.weight"
        assert dmp.state_dict()[sd_key].shape == (V2 // W, D)                   # this rank's row block
        torch.testing.assert_close(inner.embed_collection.full_state_dict()["embeddings.item_embedding.weight"], table0,
                                   rtol=0, atol=0)

        per = [torch.from_numpy(np.random.default_rng(900 + r).integers(1, V2, size=(B2, T2))) for r in range(W)]
        seqs = per[rank]
        kjt = t.KeyedJaggedTensor.from_lengths_sync(keys=["item"], values=torch.cat(list(seqs), dim=0),
                                                    lengths=torch.IntTensor([T2] * B2)).to(dev)     # torchrec/train.py:33-41
        dmp.train()
        w_out, b_out = inner.out.weight.detach().clone().cpu(), inner.out.bias.detach().clone().cpu()
        logits = dmp(kjt)
        loss = nn.CrossEntropyLoss()(logits.view(-1, V2), seqs.view(-1).to(dev))
        optimizer.zero_grad()
        loss.backward()
        optimizer.step()

        # ---- restated without sharding: gather -> linear -> CE on every rank's batch, ADAM on the union
        grads, ids_all = [], []
        for r in range(W):
            emb = table0[per[r].view(-1)].clone().requires_grad_(True)
            l_r = nn.CrossEntropyLoss()((emb @ w_out.T + b_out).view(-1, V2), per[r].view(-1))
            l_r.backward()
            grads.append(emb.grad.numpy())
            ids_all.append(per[r].view(-1).numpy())
            if r == rank:
                np.testing.assert_allclose(loss.item(), l_r.item(), rtol=1e-5)
        w_ref = table0.numpy().copy()
        m_ref, v_ref = np.zeros_like(w_ref), np.zeros_like(w_ref)
        g_idx = np.concatenate(ids_all).astype(np.int64)
        n = g_idx.size
        orc.embedding_backward_update(w_ref, D, m_ref, v_ref, np.concatenate(grads).astype(np.float32), g_idx,
                                      np.arange(0, n + 1, T2, dtype=np.int64), [0], [V2], None, 1, W * B2, orc.POOL_NONE,
                                      orc.OptArgs(orc.OPT_ADAM, 3e-2, 1e-8, 1e-4, 0.9, 0.999, 1))
        got = inner.embed_collection.full_state_dict()["embeddings.item_embedding.weight"].numpy()
        np.testing.assert_allclose(got, w_ref, rtol=1e-5, atol=2e-6)
        # DDP kept the dense half identical on both ranks
        both = [None] * W
        dist.all_gather_object(both, inner.out.weight.detach().cpu().numpy())
        np.testing.assert_array_equal(both[0], both[1])
        assert not np.array_equal(both[0], w_out.numpy())
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_dmp_sequence_collection_fused_adam_world2():
    port = free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=run_rank_sequence, args=(r, port, results)) for r in range(W)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=240)
        for p in procs:
            if p.is_alive():
                p.terminate()
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(results) == {0: "ok", 1: "ok"}
