"""Tuning helper (not product): several builds of the update kernel in ONE process, on the same sorted input.

    python tools/tune_update_multi.py base=tdfo_b200/libtdfo_b200.so st2m3=tools/_variants/st2m3.so ...

For each workload shape (BASELINE config 2; the shape one rank of 8 sees table-wise: F=3, B=131072;
config 4's row-wise shard of one rank in bf16) the first
library named sorts the batch once; then every library runs `tdfo_sparse_update` on a fresh copy of the tables
and state (compared BIT FOR BIT with the first library's result: the variants keep its arithmetic and summation
order) and is timed over `REPS` launches on rotating batches."""
import os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pathlib import Path
from tdfo_b200 import ops, _lib

REPS = int(os.environ.get("REPS", 10))
libs = [(a.split("=")[0], Path(a.split("=")[1]).resolve()) for a in sys.argv[1:]]
handles = {}


def use(name):
    """Point the ctypes binding at another build of the library (each .so is its own CUDA module)."""
    if name not in handles:
        _lib._lib = None
        _lib.LIB_PATH = dict(libs)[name]
        handles[name] = _lib.load()
    _lib._lib = handles[name]


dev = torch.device("cuda")
shapes = [("config2", 26, 1_000_000, 128, 20, 16384, torch.float32), ("1of8_tw", 3, 1_000_000, 128, 20, 131072, torch.float32),
          ("bf16_rw", 8, 6_250_000, 128, 20, 8192, torch.bfloat16),
          # config 2's batch with every id distinct within its table: no row has a second occurrence
          ("config2_uniq", 26, 1_000_000, 128, 20, 16384, torch.float32)]
if os.environ.get("SHAPES"):
    shapes = [s for s in shapes if s[0] in os.environ["SHAPES"].split(",")]
for tag, F, rows, D, P, B, dt in shapes:
    use(libs[0][0])
    gen = torch.Generator(device=dev).manual_seed(42)
    w0 = torch.empty((F * rows, D), device=dev).uniform_(-0.001, 0.001).to(dt)
    w = torch.empty_like(w0)
    state = torch.zeros(F * rows, device=dev)
    grp = ops.TableGroup(weights=w, feat_row_base=torch.arange(F, device=dev) * rows,
                         feat_rows=torch.full((F,), rows, device=dev, dtype=torch.int64),
                         feat_out_col=torch.arange(F, device=dev) * D, out_stride=F * D, state0=state)
    lengths = torch.full((F * B,), P, dtype=torch.int32, device=dev)
    offsets = ops.complete_cumsum(lengths)
    L = F * B * P
    grad = torch.randn((B, F * D), device=dev, generator=gen)
    opt = ops.make_optimizer_args(_lib.OPT_ROWWISE_ADAGRAD, 0.01, 1e-8)
    lay = ops.backward_layout(L, F * rows)
    worklist = torch.empty(ops.hot_row_worklist_entries(L), dtype=torch.int32, device=dev)
    prepared = []
    for i in range(3):
        if tag.endswith("_uniq"):
            idx = torch.cat([torch.randperm(rows, device=dev, generator=gen)[:B * P] for _ in range(F)])
        else:
            idx = torch.randint(0, rows, (L,), device=dev, generator=gen)
        k, v = ops.linearize_indices(grp, idx, offsets, B, 0)
        sk, sv = ops.radix_sort_pairs(k, v, lay.key_bits, grp.err_flag)
        uk, ss, nu, fv = ops.unique_segments(sk, grp.err_flag, sv)
        prepared.append((sv, uk, ss, fv, nu))
    U = int(prepared[0][4].item())
    esz = w0.element_size()
    upd_bytes = F * B * D * 4 + U * (2 * D * esz + 8)
    print(f"## {tag}: F={F} rows={rows} D={D} {dt} B={B} P={P} L={L} unique={U} algorithmic bytes {upd_bytes / 1e9:.3f} GB", flush=True)
    ref_w = ref_s = None
    for name, _ in libs:
        use(name)
        w.copy_(w0); state.zero_()
        sv, uk, ss, fv, nu = prepared[0]
        ops.sparse_update(grp, grad, sv, uk, ss, fv, nu, offsets, B, L, 0, opt, worklist=worklist)
        torch.cuda.synchronize()
        if ref_w is None:
            ref_w, ref_s = w.clone(), state.clone()
            same = "reference"
        else:
            same = "bit-identical" if torch.equal(w, ref_w) and torch.equal(state, ref_s) else "DIFFERS"
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(REPS)]
        for it in range(REPS + 2):
            sv, uk, ss, fv, nu = prepared[it % 3]
            e = ev[max(it - 2, 0)]
            e[0].record()
            ops.sparse_update(grp, grad, sv, uk, ss, fv, nu, offsets, B, L, 0, opt, worklist=worklist)
            e[1].record()
        torch.cuda.synchronize()
        ms = [a.elapsed_time(b) for a, b in ev]
        m = statistics.mean(ms)
        print(f"{name:10s} upd {m:.3f} ms (min {min(ms):.3f})  {upd_bytes / m / 1e6:6.0f} GB/s  {same}", flush=True)
    if os.environ.get("FWD"):          # the pooled lookup of every build on the same batch (bit-identical outputs expected)
        fwd_bytes = L * (D * esz + 8) + (F * B + 1) * 8 + F * B * D * 4
        idxs = [torch.randint(0, rows, (L,), device=dev, generator=gen) for _ in range(3)]
        out = torch.empty((B, F * D), device=dev)
        ref_o = None
        for name, _ in libs:
            use(name)
            ops.embedding_forward(grp, idxs[0], offsets, B, 0, out=out)
            torch.cuda.synchronize()
            if ref_o is None:
                ref_o, same = out.clone(), "reference"
            else:
                same = "bit-identical" if torch.equal(out, ref_o) else "DIFFERS"
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(REPS)]
            for it in range(REPS + 2):
                e = ev[max(it - 2, 0)]
                e[0].record()
                ops.embedding_forward(grp, idxs[it % 3], offsets, B, 0, out=out)
                e[1].record()
            torch.cuda.synchronize()
            ms = [a.elapsed_time(b) for a, b in ev]
            m = statistics.mean(ms)
            print(f"{name:10s} fwd {m:.3f} ms (min {min(ms):.3f})  {fwd_bytes / m / 1e6:6.0f} GB/s  {same}", flush=True)
        del idxs, out, ref_o
    assert int(grp.err_flag.item()) == 0
    del w0, w, state, grp, grad, prepared, ref_w, ref_s, worklist
    torch.cuda.empty_cache()
