#!/usr/bin/env python
"""bench.py — embedding lookups/sec (fwd + bwd + fused update) on BASELINE config 2.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the hot path over one synthetic jagged batch:
pooled lookup -> (synthetic upstream gradient) -> linearise -> radix sort -> unique ->
fused row-wise Adagrad update.  Prints ONE JSON line (rank 0).

* `value`     device-resident: ids / offsets / gradient already in HBM, timed with CUDA events
              around exactly K steps on the launching stream, max over ranks.
* `e2e`       the same metric through the public module API (KeyedJaggedTensor ->
              EmbeddingBagCollection -> loss.backward()) with HOST (pinned) inputs: the per-step
              host->device copy and a device->host read of the loss are inside the timed region.
* `roofline`  the dominant kernel's algorithmic bytes / its CUDA-event time inside the timed region,
              against the measured HBM copy bandwidth in MEASURED_PEAKS.json.
* `cpu_baseline` / `--impl reference`: the CPU oracle (oracle/tdfo_oracle.c, OpenMP over all host
              threads) on a bounded sample of the same workload.  The reference's own CPU kernels
              (torchrec / fbgemm_gpu) are not installable here, hence kind = "port".
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "embedding lookups/sec (fwd+bwd+update)"
UNIT = "lookups/s"

# BASELINE.json configs[1]
CFG = dict(tables=26, rows=1_000_000, dim=128, pooling=20, batch=16384, lr=0.01, eps=1e-8)
WORKLOAD = ("single-GPU EmbeddingBagCollection: 26 tables x 1M rows x dim 128 fp32, pooling factor 20, "
            "batch 16384, uniform ids, SUM pooling, row-wise Adagrad")
N_BATCHES = 4      # distinct id batches rotated through the timed region


def measured_peak_gbs():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def algorithmic_bytes(L, bags, U, D, w=4, g=4, s=4):
    """SURVEY.md §8(d) / BASELINE.md §3."""
    fwd = L * (D * w + 8) + (bags + 1) * 8 + bags * D * g
    upd_kernel = bags * D * g + U * (2 * D * w + 2 * s)      # the update kernel's own share
    bwd = upd_kernel + L * 8                                  # + ids read by the linearise stage
    return fwd, bwd, upd_kernel


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index, self.proc, self.path = gpu_index, None, f"/tmp/tdfo_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.fh = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.gpu_index)], stdout=self.fh,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.fh.close()
        sm, smax, reasons = [], [], set()
        for line in open(self.path):
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); smax.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.remove(self.path)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smax), "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU arm (oracle port)
def cpu_arm(steps: int, warmup: int, sample_batch: int, rows: int | None = None):
    """Times oracle fwd + bwd/update (row-wise Adagrad) on the host.  Returns (lookups/s, info)."""
    import oracle as orc
    import psutil
    F, D, P = CFG["tables"], CFG["dim"], CFG["pooling"]
    rows = rows or CFG["rows"]
    need = F * rows * D * 4
    avail = psutil.virtual_memory().available
    if need > 0.5 * avail:                               # keep tables <= half of free host RAM
        rows = max(1000, int(0.5 * avail / (F * D * 4)))
    B = sample_batch
    # all the host threads this process may use, whatever OMP_NUM_THREADS the launcher exported
    # (torchrun sets it to 1); `cores` reports the OpenMP thread count actually in effect
    avail_cpus = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    cores = orc.set_num_threads(avail_cpus)
    w = np.empty((F * rows, D), dtype=np.float32)
    w.fill(0.01)
    h = np.zeros(F * rows, dtype=np.float32)
    rng = np.random.default_rng(42)
    lengths = np.full(F * B, P, dtype=np.int32)
    offsets = orc.complete_cumsum(lengths)
    L = int(offsets[-1])
    row_base = (np.arange(F) * rows).astype(np.int64)
    frows = np.full(F, rows, dtype=np.int64)
    col = (np.arange(F) * D).astype(np.int64)
    grad = rng.normal(size=(B, F * D)).astype(np.float32)
    opt = orc.OptArgs(orc.OPT_ROWWISE_ADAGRAD, CFG["lr"], CFG["eps"])
    times = []
    for i in range(warmup + steps):
        idx = rng.integers(0, rows, size=L).astype(np.int64)
        t0 = time.perf_counter()
        orc.embedding_forward(w, D, idx, offsets, None, row_base, frows, col, F, B, orc.POOL_SUM)
        orc.embedding_backward_update(w, D, h, None, grad, idx, offsets, row_base, frows, col, F, B, orc.POOL_SUM, opt)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    t = sum(times)
    frac = "the whole workload batch" if B == CFG["batch"] else f"1/{CFG['batch'] // B} of the workload batch"
    info = {"cores": cores, "kind": "port",
            "sample": f"{steps} steps of batch {B} ({frac}, L={L} lookups/step), "
                      f"{F} tables x {rows} rows x dim {D} fp32, oracle/tdfo_oracle.c; every stage (lookup, linearise, "
                      f"counting sort, run-length, update) runs on all {cores} OpenMP threads"}
    return L * steps / t, t / steps * 1e3, info


# ----------------------------------------------------------------------------- the reference's literal CPU mode
DDP_CFG1 = dict(tables=[("user", 6040, ["user"], 1), ("movie", 3952, ["movie", "history"], 20), ("gender", 2, ["gender"], 1),
                        ("age", 7, ["age"], 1), ("occupation", 21, ["occupation"], 1), ("zip", 3439, ["zip"], 1),
                        ("genre", 18, ["genres"], 6)], D=16, B=16, W=2)


def _ddp_rank(rank, port, steps, warmup, q):
    """One rank of BASELINE config 1 the way the reference runs on CPU (torchrec/train.py:192-195,255-260):
    plain torch modules under DistributedDataParallel over gloo, dense [V, D] embedding gradients all-reduced
    every step, a dense optimizer over every row.  None of this repository's code is on this path."""
    import torch
    import torch.distributed as dist
    import torch.nn as nn
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=DDP_CFG1["W"])
    torch.set_num_threads(max(1, (os.cpu_count() or 2) // DDP_CFG1["W"]))
    torch.manual_seed(7)
    D, B = DDP_CFG1["D"], DDP_CFG1["B"]
    feats = [(f, rows, mx if f not in ("movie",) else 1) for _, rows, fs, mx in DDP_CFG1["tables"] for f in fs]

    class Model(nn.Module):
        def __init__(self):
            super().__init__()
            self.bags = nn.ModuleDict({name: nn.EmbeddingBag(rows, D, mode="sum") for name, rows, _, _ in DDP_CFG1["tables"]})
            self.mlp = nn.Sequential(nn.Linear(len(feats) * D, 64), nn.ReLU(), nn.Linear(64, 32), nn.ReLU(), nn.Linear(32, 1))
            self.table_of = {f: name for name, _, fs, _ in DDP_CFG1["tables"] for f in fs}

        def forward(self, batch):
            return self.mlp(torch.cat([self.bags[self.table_of[f]](ids, off) for f, (ids, off) in batch.items()], dim=1))

    model = nn.parallel.DistributedDataParallel(Model())
    opt = torch.optim.Adam(model.parameters(), lr=3e-4, weight_decay=1e-4)       # torchrec/train.py:258-260
    bce = nn.BCEWithLogitsLoss()
    rng = np.random.default_rng(100 + rank)
    lookups = 0
    t0 = None
    for i in range(warmup + steps):
        if i == warmup:
            dist.barrier()
            t0 = time.perf_counter()
        batch = {}
        for f, rows, mx in feats:
            ln = np.ones(B, np.int64) if mx == 1 else rng.integers(1, mx + 1, size=B)
            off = np.concatenate([[0], np.cumsum(ln)[:-1]])
            batch[f] = (torch.from_numpy(rng.integers(0, rows, size=int(ln.sum()))), torch.from_numpy(off))
            if i >= warmup:
                lookups += int(ln.sum())
        loss = bce(model(batch), torch.from_numpy(rng.integers(0, 2, size=(B, 1)).astype(np.float32)))
        opt.zero_grad()
        loss.backward()
        opt.step()
    dist.barrier()
    dt = time.perf_counter() - t0
    tot = torch.tensor([float(lookups)])
    dist.all_reduce(tot)
    if rank == 0:
        q.put((float(tot.item()) / dt, dt / steps * 1e3))
    dist.destroy_process_group()


def cpu_ddp_baseline(steps=200, warmup=20):
    """BASELINE config 1 timed in the reference's own CPU mode; returns the `cpu_baseline_ddp` block."""
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s_:
        s_.bind(("127.0.0.1", 0))
        port = s_.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ddp_rank, args=(r, port, steps, warmup, q)) for r in range(DDP_CFG1["W"])]
    for p_ in procs:
        p_.start()
    value = ms = None
    deadline = time.time() + 300
    while time.time() < deadline:
        try:
            value, ms = q.get(timeout=1.0)
            break
        except Exception:
            if not any(p_.is_alive() for p_ in procs):       # a rank died: report the block as unmeasured
                break
    for p_ in procs:
        p_.join(timeout=60)
        if p_.is_alive():
            p_.terminate()
    return {"value": value, "unit": UNIT, "ms_per_step": ms, "world_size": DDP_CFG1["W"], "steps": steps,
            "kind": "reference mode restated with torch modules (torchrec itself is not installable here)",
            "workload": "config 1: MovieLens-1M-shaped EmbeddingBag tables (user 6040, movie 3952 + history bag, gender, age, "
                        "occupation, zip, genres) dim 16 + MLP(64,32,1), batch 16 per rank, DistributedDataParallel over gloo, "
                        "dense embedding gradients all-reduced, dense Adam over every row (torchrec/train.py:255-260)"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # same workload, same batch, same step / warm-up counts as the GPU arm (about one second per step on 16 cores)
    value, ms, info = cpu_arm(args.steps, args.warmup, sample_batch=CFG["batch"])
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU arm: oracle port of the path on the host cores; the reference's own "
                   "CPU kernels (torchrec/fbgemm_gpu) are not installable on this image"},
        "cpu_baseline": {"value": value, "unit": UNIT, **info},
        "cpu_baseline_ddp": cpu_ddp_baseline(),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    import tdfo_b200 as t
    from tdfo_b200 import _lib, ops

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the tdfo_b200 path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)

    F, rows, D, P, B = CFG["tables"], CFG["rows"], CFG["dim"], CFG["pooling"], CFG["batch"]
    gen = torch.Generator(device=dev).manual_seed(42 + rank)
    bound = 1.0 / (rows ** 0.5)

    tables = [t.EmbeddingBagConfig(num_embeddings=rows, embedding_dim=D, name=f"t{i}", feature_names=[f"f{i}"],
                                   pooling=t.PoolingType.SUM) for i in range(F)]
    ebc = t.EmbeddingBagCollection(tables, device=dev)
    fused = ebc.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=CFG["lr"], eps=CFG["eps"])
    ebc.train()
    g0 = ebc._groups[0]
    grp = ebc._table_group(g0, True)
    keys = [f"f{i}" for i in range(F)]

    lengths = torch.full((F * B,), P, dtype=torch.int32, device=dev)
    offsets = ops.complete_cumsum(lengths)
    L = int(offsets[-1].item())
    bags = F * B
    row_base = torch.arange(F, device=dev).repeat_interleave(B * P) * rows
    batches, uniques = [], []
    for bi in range(N_BATCHES):
        if args.ids == "zipf":        # hot rows: ids ~ Zipf(1.05) clipped to the table
            z = np.random.default_rng(42 + bi).zipf(1.05, size=L)
            idx = torch.from_numpy(np.minimum(z - 1, rows - 1).astype(np.int64)).to(dev)
        else:
            idx = torch.randint(0, rows, (L,), device=dev, generator=gen)
        batches.append(idx)
        uniques.append(int(torch.unique(idx + row_base).numel()))
    del row_base
    grad_out = torch.randn((B, F * D), device=dev, generator=gen)
    opt = ops.make_optimizer_args(_lib.OPT_ROWWISE_ADAGRAD, CFG["lr"], CFG["eps"])
    lay = ops.backward_layout(L, F * rows)
    scratch = torch.empty(lay.total_bytes, dtype=torch.uint8, device=dev)
    out = torch.empty((B, F * D), dtype=torch.float32, device=dev)
    stages = ["forward", "linearize", "sort", "unique", "update"]
    worklist = torch.empty(ops.hot_row_worklist_entries(L), dtype=torch.int32, device=dev)      # hot-row queue of the staged update

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_fused(i):
        idx = batches[i % N_BATCHES]
        # the forward also emits the (row, bag) pairs the backward sorts (linearise stage fused away)
        ops.embedding_forward(grp, idx, offsets, B, _lib.POOL_SUM, out=out, backward_scratch=scratch)
        ops.embedding_backward_update(grp, grad_out, idx, offsets, B, _lib.POOL_SUM, opt, scratch=scratch,
                                      keys_in_scratch=True)

    def step_staged(i, ev):
        """Same kernels, same order, one C-ABI call per stage so each can be bracketed by events."""
        idx = batches[i % N_BATCHES]
        ev[0].record()
        ops.embedding_forward(grp, idx, offsets, B, _lib.POOL_SUM, out=out)
        ev[1].record()
        k, v = ops.linearize_indices(grp, idx, offsets, B, _lib.POOL_SUM)
        ev[2].record()
        sk, sv = ops.radix_sort_pairs(k, v, lay.key_bits, grp.err_flag)
        ev[3].record()
        uk, ss, nu, fv = ops.unique_segments(sk, grp.err_flag, sv)
        ev[4].record()
        ops.sparse_update(grp, grad_out, sv, uk, ss, fv, nu, offsets, B, L, _lib.POOL_SUM, opt, worklist=worklist)
        ev[5].record()

    # ---------------- device-resident arm (`value`): the single-call product path
    for i in range(args.warmup):
        step_fused(i)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = ops.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step_fused(args.warmup + i)
    e1.record()
    barrier()
    launches = ops.launch_count() - launches0
    ms_total = e0.elapsed_time(e1)

    # ---------------- per-kernel timing with the same kernels staged (roofline of the dominant kernel)
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(args.steps)]
    for i in range(2):
        step_staged(i, evs[0])
    barrier()
    for i in range(args.steps):
        step_staged(args.warmup + i, evs[i])
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    stage_ms = {s: statistics.mean(evs[i][j].elapsed_time(evs[i][j + 1]) for i in range(args.steps))
                for j, s in enumerate(stages)}
    ops.check_error_flag(grp.err_flag, "bench")

    if world > 1:
        tt = torch.tensor([ms_total], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total = float(tt.item())
    ms_step = ms_total / args.steps
    value = L * world * args.steps / (ms_total * 1e-3)

    U = statistics.mean(uniques)
    fwd_b, bwd_b, upd_b = algorithmic_bytes(L, bags, U, D)
    peak, peak_src = measured_peak_gbs()
    kern_bytes = {"forward": fwd_b, "update": upd_b}
    dominant = max(("forward", "update"), key=lambda s: stage_ms[s])
    achieved = kern_bytes[dominant] / (stage_ms[dominant] * 1e-3) / 1e9
    traffic = None
    tp = ROOT / "profiles" / "ncu_traffic.json"
    if tp.exists():
        try:
            traffic = json.loads(tp.read_text()).get(dominant)
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": {"forward": "pooled_forward_kernel", "update": "sparse_update_kernel"}[dominant],
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": kern_bytes[dominant],
                "launch_ms": stage_ms[dominant]}
    step_gbs = (fwd_b + bwd_b) / (ms_step * 1e-3) / 1e9
    kernels = {s: {"ms": stage_ms[s]} for s in stages}
    for s in ("forward", "update"):
        kernels[s]["GB/s"] = kern_bytes[s] / (stage_ms[s] * 1e-3) / 1e9
        kernels[s]["frac_of_peak"] = kernels[s]["GB/s"] / peak

    # ---------------- end-to-end arm through the public module API with host buffers
    e2e = run_e2e(args, t, ebc, keys, lengths, batches, grad_out, F, B, D, L, world, dev, barrier)

    # ---------------- CPU baseline on rank 0 at N=1
    cpu = cpu_ddp = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        del ebc, grp, scratch, out, batches
        torch.cuda.empty_cache()
        v, _, info = cpu_arm(steps=3, warmup=1, sample_batch=B)
        cpu = {"value": v, "unit": UNIT, **info}
        cpu_ddp = cpu_ddp_baseline()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD if args.ids == "uniform" else WORKLOAD.replace("uniform ids", "Zipf(1.05) ids"),
                       "lookups_per_step_per_gpu": L, "unique_rows_per_step": U,
                       "l2": f"inputs exceed L2: 13.3 GB of tables, {N_BATCHES} rotating id batches of {L * 8 >> 20} MiB",
                       "parallelism": "1 GPU" if world == 1 else f"{world} independent table shards (no collective)",
                       "optimizer": "row-wise Adagrad lr 0.01 eps 1e-8, fused into backward"},
            "roofline": roofline,
            "step_roofline": {"algorithmic_GB_per_step": (fwd_b + bwd_b) / 1e9, "achieved_GBps": step_gbs,
                              "frac_of_measured_peak": step_gbs / peak, "frac_of_8TBps_spec": step_gbs / 8000.0},
            "kernels": kernels,
            "cpu_baseline": cpu,
            "cpu_baseline_ddp": cpu_ddp,
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _measured_traffic(torch, dist, m, sp, world, rank, B, dev, lengths, ids):
    """Exact per-rank counts for one batch (outside the timed region): ids received per local slot, unique
    rows per slot (torch.unique on what landed), and the bytes this rank sends over NVLink.  Returns the
    algorithmic HBM bytes of the rank (SURVEY.md §8d formula, per slot) and the NVLink send bytes."""
    D, w = sp["D"], (2 if sp["bf16"] else 4)
    W, r = world, rank
    GB = W * B
    with torch.no_grad():
        c = m._input_dist(ids, lengths, B)                 # collective: the exchange alone, nothing consumes it
    torch.cuda.synchronize()
    z = m._peer.zones[c["slot"]]
    Fl = m._F_local
    off = z.off_x[: Fl * GB + 1]
    bounds = off[::GB].tolist() if Fl else [0]
    widths = [D] * m._n_tw[r] + [m._Dc] * m._n_cw[r] + [D] * m._n_rw
    hbm, U_total = 0.0, 0
    for j in range(Fl):
        Lj = bounds[j + 1] - bounds[j]
        Uj = int(torch.unique(z.ids_x[bounds[j]:bounds[j + 1]]).numel()) if Lj else 0
        Dj = widths[j]
        hbm += Lj * (Dj * w + 8) + (GB + 1) * 8 + GB * Dj * 4            # lookup
        hbm += GB * Dj * 4 + Lj * 8 + Uj * (2 * Dj * w + 8)              # backward + row-wise Adagrad update
        U_total += Uj
    L_recv = bounds[-1]
    # NVLink send bytes of this rank
    off_local = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(lengths.long(), 0)])
    per_f = (off_local[B::B] - off_local[:-1:B]).tolist()                # my ids per feature
    send = 0
    for d in range(W):
        if d == r:
            continue
        for f in m._tw_feats[d] + [f_ for (f_, _) in m._cw_slots[d]]:
            send += per_f[f] * 8 + B * 4                                 # ids + lengths to the owner
        send += len(m._tw_feats[d]) * B * D * 4 + len(m._cw_slots[d]) * B * m._Dc * 4     # my gradient rows to the owner
    send += (W - 1) * (m._n_tw[r] * B * D * 4 + m._n_cw[r] * B * m._Dc * 4)              # pooled rows I return
    if m._n_rw:
        for k_, f in enumerate(m._rw_feats):
            seg = ids[off_local[f * B]:off_local[(f + 1) * B]]
            send += int(((seg // m._rw_block[k_]) != r).sum().item()) * 8 + (W - 1) * B * 4
        send += 2 * (W - 1) * B * m._n_rw * D * 4                        # partial sums out, gradient rows out
    stats = torch.tensor([hbm, float(send), float(L_recv), float(U_total)], dtype=torch.float64, device=dev)
    mx = stats.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = stats.clone()
    dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    return {"hbm_bytes_max_rank": float(mx[0]), "hbm_bytes_mean_rank": float(sm[0]) / W,
            "nvlink_send_bytes_max_rank": float(mx[1]), "nvlink_send_bytes_mean_rank": float(sm[1]) / W,
            "ids_received_max_rank": float(mx[2]), "ids_received_mean_rank": float(sm[2]) / W,
            "unique_rows_max_rank": float(mx[3]), "unique_rows_mean_rank": float(sm[3]) / W}


def sharded_workload(cfg, args, rank, world, dev, headline):
    """Parity check + timed run of BASELINE config `cfg` sharded over the `world` GPUs (weak scaling in the
    batch, except config 5 whose global batch is fixed at 65536)."""
    import gc
    import torch
    import torch.distributed as dist
    import tdfo_b200 as t
    import bench_workloads as bw
    from tdfo_b200 import distributed as td, ops

    sp = bw.spec(cfg, world)
    plan_override = args.plan if (cfg == 2 and args.plan != "cw") else None
    parity = bw.parity_check(cfg, world, rank, dev, plan_override=plan_override)

    rows, D, B = sp["rows"], sp["D"], sp["B"]
    F = len(rows)
    tables = bw.make_tables(t, sp)
    plan = bw.make_plan(td, sp, tables, world, plan_override)
    kinds = [plan[tb.name].sharding_type.value for tb in tables]
    n_batches = N_BATCHES if cfg == 2 else 2
    batches = [bw.device_batch(torch, sp, rows, B, dev, seed=42 + rank + 1000 * bi) for bi in range(n_batches)]
    Ls = [int(b[1].numel()) for b in batches]
    tot = torch.tensor([float(sum(Ls)) / n_batches, float(max(Ls))], dtype=torch.float64, device=dev)
    dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    L_global, cap = float(tot[0]), int(tot[1]) + 1024      # no rank can receive more ids than exist
    m = td.ShardedEmbeddingBagCollection(tables, plan, dev, init_on_device=True, max_ids_per_rank=cap)
    m.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=CFG["lr"], eps=CFG["eps"])
    m.train()
    keys = [f"f{i}" for i in range(F)]
    gen = torch.Generator(device=dev).manual_seed(42 + rank)
    grad_out = torch.randn((B, F * D), device=dev, generator=gen)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    prefetch = args.prefetch_input_dist

    def make_kjt(i):
        lengths, ids = batches[i % n_batches]
        return t.KeyedJaggedTensor(keys, ids, lengths=lengths, stride=B)

    state = {"next": None}

    def step(i, last):
        kjt = state["next"] if state["next"] is not None else make_kjt(i)
        state["next"] = None
        out = m(kjt).values()
        if prefetch and not last:
            state["next"] = make_kjt(i + 1)
            m.prefetch_input_dist(state["next"])           # batch i+1's exchange runs next to this backward
        out.backward(grad_out)

    for i in range(args.warmup):
        step(i, i + 1 == args.warmup)
    barrier()
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0"))) if headline else None
    if sampler is not None and rank == 0:
        sampler.start()
    launches0 = ops.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i, i + 1 == args.steps)
    e1.record()
    barrier()
    launches = ops.launch_count() - launches0
    tt = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_step = float(tt.item()) / args.steps
    value = L_global / (ms_step * 1e-3)
    m.check_errors()

    e2e = None
    if headline:
        lengths0 = batches[0][0]
        same_len = all(torch.equal(b[0], lengths0) for b in batches)
        if same_len:
            e2e = run_e2e(args, t, m, keys, lengths0, [b[1] for b in batches], grad_out, F, B, D, Ls[0], world, dev,
                          barrier)
    clocks = sampler.stop() if (sampler is not None and rank == 0) else None
    traffic = _measured_traffic(torch, dist, m, sp, world, rank, B, dev, *batches[0])
    peak, peak_src = measured_peak_gbs()
    hbm_gbs = traffic["hbm_bytes_max_rank"] / (ms_step * 1e-3) / 1e9
    nv_gbs = traffic["nvlink_send_bytes_max_rank"] / (ms_step * 1e-3) / 1e9
    n_cw = kinds.count("column_wise")
    res = {
        "workload": sp["name"], "value": value, "unit": UNIT, "ms_per_step": ms_step,
        "lookups_per_step_global": L_global, "batch_per_gpu": B, "dtype": "bf16" if sp["bf16"] else "f32",
        "plan": f"{kinds.count('table_wise')} table-wise + {n_cw} column-wise + {kinds.count('row_wise')} row-wise over "
                f"{world} GPUs; every exchange (lengths, ids, pooled rows / partial sums, gradient rows) is written by "
                "the kernels into NVLink peer memory, no collective library call on the step",
        "prefetch_input_dist": bool(prefetch),
        "hbm": {"algorithmic_bytes_per_step_max_rank": traffic["hbm_bytes_max_rank"],
                "algorithmic_bytes_per_step_mean_rank": traffic["hbm_bytes_mean_rank"],
                "achieved_GBps": hbm_gbs, "peak": peak, "frac": hbm_gbs / peak, "frac_of_8TBps_spec": hbm_gbs / 8000.0,
                "unique_rows": "measured (torch.unique on the ids each rank received)",
                "unique_rows_mean_rank": traffic["unique_rows_mean_rank"],
                "ids_received_max_rank": traffic["ids_received_max_rank"],
                "ids_received_mean_rank": traffic["ids_received_mean_rank"]},
        "nvlink": {"send_bytes_per_gpu_per_step_max_rank": traffic["nvlink_send_bytes_max_rank"],
                   "send_bytes_per_gpu_per_step_mean_rank": traffic["nvlink_send_bytes_mean_rank"],
                   "GBps_if_serial": nv_gbs, "peak_GBps_per_direction": 900.0, "frac": nv_gbs / 900.0},
        "parity_check": parity, "gpu_launches": int(launches),
    }
    extra = {"e2e": e2e, "clocks": clocks, "peak_src": peak_src, "kinds": kinds, "L_local": Ls[0]}
    del m, batches, grad_out
    gc.collect()
    torch.cuda.empty_cache()
    return res, extra


def sharded_ec_block(args, rank, world, dev):
    """The module the reference itself shards (torchrec/models.py:150-164, torchrec/train.py:241-247): one item table in
    an EmbeddingCollection (one row per id), row-wise over the GPUs, fused ADAM (torchrec/train.py:191,236-240), at the
    reference's own shape and at a large one.  fwd + bwd + update per step, ids resident on the device."""
    import gc
    import torch
    import torch.distributed as dist
    import tdfo_b200 as t
    from tdfo_b200 import distributed as td
    out = {}
    shapes = {"reference shape: batch 16 per GPU, 20 ids per sequence, dim 16": (16, 20, 16, 100_000),
              "batch 16384 per GPU, 20 ids per sequence, dim 128": (16384, 20, 128, 1_000_000)}
    for name, (B, T, D, V) in shapes.items():
        tables = [t.EmbeddingConfig(name="item_embedding", embedding_dim=D, num_embeddings=V, feature_names=["item"],
                                    weight_init_max=1.0, weight_init_min=-1.0)]
        m = td.ShardedEmbeddingCollection(tables, {"item_embedding": td.row_wise()}, dev, init_on_device=True,
                                          max_ids_per_rank=world * B * T + 1024, max_local_ids=B * T)
        m.set_fused_optimizer(t.EmbOptimType.ADAM, learning_rate=3e-4, weight_decay=1e-4)
        m.train()
        gen = torch.Generator(device=dev).manual_seed(7 + rank)
        ids = [torch.randint(1, V, (B * T,), device=dev, generator=gen) for _ in range(2)]
        lengths = torch.full((B,), T, dtype=torch.int32, device=dev)
        grad = torch.randn((B * T, D), device=dev, generator=gen)

        def step(i):
            jt = m(t.KeyedJaggedTensor(["item"], ids[i % 2], lengths=lengths, stride=B))["item"]
            jt.values().backward(grad)

        for i in range(args.warmup):
            step(i)
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            step(i)
        e1.record()
        dist.barrier()
        torch.cuda.synchronize()
        tt = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item()) / args.steps
        m.check_errors()
        out[name] = {"ms_per_step": ms, "value": world * B * T / (ms * 1e-3), "unit": UNIT, "rows": V,
                     "exchange": "NVLink peer memory" if m._use_peer else "NCCL all-to-all"}
        del m, ids, grad
        gc.collect()
        torch.cuda.empty_cache()
    return out


def run_ours_sharded(args):
    """N > 1.  Headline: BASELINE config 2 (or --config) with its tables sharded over the N GPUs of the box,
    every rank feeding its own batch.  Each workload is preceded by a seeded value check against the oracle
    (`parity_check`).  On 8 GPUs configs 3, 4 and 5 run after the headline and are reported under `configs`."""
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)

    res, extra = sharded_workload(args.config, args, rank, world, dev, headline=True)
    others = {}
    want = [c for c in (args.configs.split(",") if args.configs not in ("auto", "none") else
                        (["3", "4", "5"] if (args.configs == "auto" and world == 8 and args.config == 2) else []))]
    for c in want:
        try:
            r_, _ = sharded_workload(int(c), args, rank, world, dev, headline=False)
            others[c] = r_
        except Exception as e:                              # keep the headline even if a side workload fails
            others[c] = {"error": f"{type(e).__name__}: {e}"[:400]}
    ec = None
    if args.config == 2:
        try:
            ec = sharded_ec_block(args, rank, world, dev)
        except Exception as e:
            ec = {"error": f"{type(e).__name__}: {e}"[:400]}
    ok = res["parity_check"]["ok"] and all(o.get("parity_check", {}).get("ok", False) for o in others.values()
                                            if "error" not in o)
    if rank == 0:
        kinds = extra["kinds"]
        line = {
            "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
            "scaling": "strong" if args.config == 5 else "weak", "vs_baseline": None, "dtype": res["dtype"],
            "data": "synthetic",
            "config": {"workload": res["workload"], "lookups_per_step_per_gpu": extra["L_local"],
                       "l2": "inputs exceed L2: GBs of tables per GPU, rotating id batches",
                       "parallelism": res["plan"] + ("; a column shard keeps its own row-wise Adagrad state, as "
                                                     "torchrec's does" if "column_wise" in kinds else ""),
                       "optimizer": "row-wise Adagrad lr 0.01 eps 1e-8, fused into backward",
                       "prefetch_input_dist": res["prefetch_input_dist"]},
            "roofline": {"bound": "hbm", "kernel": "whole sharded step per GPU (lookup + sort/unique + update + exchanges), "
                                                   "the rank with the most bytes",
                         "achieved": res["hbm"]["achieved_GBps"], "peak": res["hbm"]["peak"], "unit": "GB/s",
                         "frac": res["hbm"]["frac"], "traffic": None, "peak_source": extra["peak_src"],
                         "algorithmic_bytes_per_launch": res["hbm"]["algorithmic_bytes_per_step_max_rank"],
                         "launch_ms": res["ms_per_step"],
                         "note": "per-kernel rooflines are measured at N=1; unique rows per GPU are measured, not estimated",
                         "detail": res["hbm"]},
            "nvlink": res["nvlink"],
            "parity_check": res["parity_check"],
            "cpu_baseline": None,
            "e2e": extra["e2e"],
            "gpu_launches": res["gpu_launches"],
            "clocks": extra["clocks"],
        }
        if others:
            line["configs"] = others
        if ec is not None:
            line["sharded_embedding_collection"] = ec
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if not ok:
        raise SystemExit("bench.py: parity_check failed (see the JSON line)")


def run_e2e(args, t, ebc, keys, lengths_dev, batches, grad_out, F, B, D, L, world, dev, barrier):
    """Public API, host inputs: pinned ids+lengths -> H2D (prefetched two steps ahead on a copy stream,
    as the reference's DataLoader(pin_memory)/prefetch_to_device does) -> KJT -> EBC -> loss ->
    backward (fused update) -> loss.item()."""
    import torch
    import torch.distributed as dist
    # DEPTH batches are on their way to the device while one is being consumed: the reference's own
    # idiom is prefetch_to_device(size=2) (jax-flax/train_dp.py:211).  With 8 GPUs copying at once a
    # 68 MB batch takes 1.9-2.9 ms (tools/h2d_contention.py), close to the 3.3 ms step: one step of
    # look-ahead leaves no slack for jitter, two do.
    DEPTH = 2
    # ids cross PCIe as int32 (every table has far fewer than 2^31 rows) and are widened on the device by the
    # KeyedJaggedTensor (tdfo_kjt_widen_ids): half the bytes of the copy that bounds the 8-GPU end-to-end rate
    host_vals = [b.cpu().to(torch.int32).pin_memory() for b in batches]
    host_lens = lengths_dev.cpu().pin_memory()
    copy_stream = torch.cuda.Stream(device=dev)
    slots = [dict(v=torch.empty(L, dtype=torch.int32, device=dev), l=torch.empty(F * B, dtype=torch.int32, device=dev),
                  ev=torch.cuda.Event()) for _ in range(DEPTH + 1)]
    dist_prefetch = bool(getattr(args, "prefetch_input_dist", False)) and hasattr(ebc, "prefetch_input_dist") and world > 1
    pending = {}
    main = torch.cuda.current_stream(dev)

    nocopy = os.environ.get("TDFO_E2E_NOCOPY") == "1"      # diagnostic only: marks the result as not end-to-end
    if nocopy:
        for k_, s_ in enumerate(slots):
            s_["v"].copy_(batches[k_ % len(batches)])
            s_["l"].copy_(lengths_dev)

    def prefetch(i):
        s = slots[i % (DEPTH + 1)]
        copy_stream.wait_stream(main)                       # slot reuse: previous consumer is done
        if nocopy:
            s["ev"].record(copy_stream)
            return
        with torch.cuda.stream(copy_stream):
            s["v"].copy_(host_vals[i % len(host_vals)], non_blocking=True)
            s["l"].copy_(host_lens, non_blocking=True)
            s["ev"].record(copy_stream)

    loss_host = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]
    grad_flat = grad_out.view(-1)
    losses = []

    def step(i, n_total):
        s = slots[i % (DEPTH + 1)]
        main.wait_event(s["ev"])
        if i + DEPTH < n_total:
            prefetch(i + DEPTH)
        kjt = pending.pop(i, None)
        if kjt is None:
            kjt = t.KeyedJaggedTensor.from_lengths_sync(keys, s["v"], s["l"])
        kt = ebc(kjt)
        if dist_prefetch and i + 1 < n_total:
            # batch i+1 is already on the device (copied two steps ahead): start its lengths / ids exchange now
            s1 = slots[(i + 1) % (DEPTH + 1)]
            main.wait_event(s1["ev"])
            pending[i + 1] = t.KeyedJaggedTensor.from_lengths_sync(keys, s1["v"], s1["l"])
            ebc.prefetch_input_dist(pending[i + 1])
        loss = torch.dot(kt.values().view(-1), grad_flat)     # scalar loss; d(loss)/d(pooled) = grad_out
        loss.backward()
        # device->host read of the step's loss: asynchronous copy into pinned memory, consumed one step
        # later (logging lags by a step, the GPU never waits for the host)
        loss_host[i % 2].copy_(loss.detach().view(1), non_blocking=True)
        loss_ev[i % 2].record(main)
        if i > 0:
            loss_ev[(i - 1) % 2].synchronize()
            losses.append(float(loss_host[(i - 1) % 2][0]))
        if i + 1 == n_total:
            loss_ev[i % 2].synchronize()
            losses.append(float(loss_host[i % 2][0]))

    n_w = max(args.warmup, 1)
    for j in range(min(DEPTH, n_w)):
        prefetch(j)
    for i in range(n_w):
        step(i, n_w)
    barrier()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for j in range(min(DEPTH, args.steps)):
        prefetch(j)
    for i in range(args.steps):
        step(i, args.steps)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    wall_ms = (time.perf_counter() - t0) * 1e3
    ms = max(ms, 0.0)
    if world > 1:
        tt = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    return {"value": L * world * args.steps / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / args.steps,
            "wall_ms_per_step": wall_ms / args.steps,
            "h2d_bytes_per_step": 0 if nocopy else L * 4 + F * B * 4, "d2h_bytes_per_step": 4,
            "id_format": "int32 on the host and over PCIe, widened to int64 on the device",
            "prefetch_input_dist": dist_prefetch,
            "losses_read": len(losses),
            "prefetch_depth": DEPTH,
            "api": "pinned int32 ids / lengths -> H2D -> KeyedJaggedTensor.from_lengths_sync -> (Sharded)EmbeddingBagCollection"
                   ".forward -> loss.backward() (fused update) -> loss D2H"}


def main():
    # keep stdout to the one JSON line: NCCL prints its version banner there at NCCL_DEBUG=VERSION
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ids", choices=["uniform", "zipf"], default="uniform",
                    help="N=1: id distribution; zipf = Zipf(1.05) hot-row variant of config 2 (reported separately)")
    ap.add_argument("--config", type=int, choices=[2, 3, 4, 5], default=2,
                    help="BASELINE config run as the headline (3-5 need N > 1; default 2)")
    ap.add_argument("--configs", default="auto",
                    help="N>1: further configs reported under `configs` after the headline: auto (3,4,5 on 8 GPUs), "
                         "none, or a comma list such as 3,4")
    ap.add_argument("--no-prefetch-input-dist", dest="prefetch_input_dist", action="store_false",
                    help="N>1: do NOT issue batch i+1's lengths / ids exchange before batch i's backward (it is on by "
                         "default: 3.00 vs 3.17 ms/step on 8 GPUs, profiles/r2_bench_n8*.json)")
    ap.add_argument("--plan", choices=["tw", "mixed", "cw"], default="cw",
                    help="N>1: cw = table-wise round-robin with the tables that do not divide evenly split column-wise "
                         "(balanced; 21.1 G lookups/s at N=8), tw = all table-wise (19.5), mixed = remainder row-wise (16.7)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        if args.warmup < 3:
            args.warmup = 3
        if int(os.environ.get("WORLD_SIZE", "1")) > 1:
            run_ours_sharded(args)
        elif args.config != 2:
            raise SystemExit("bench.py: configs 3-5 are sharded workloads, run them with --gpus >= 2 under torchrun")
        else:
            run_ours(args)


if __name__ == "__main__":
    main()
