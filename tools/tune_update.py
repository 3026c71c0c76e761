"""Tuning helper (not product): times the staged kernels on BASELINE config 2 for each TDFO_*_VARIANT."""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pathlib import Path
from tdfo_b200 import ops, _lib
if os.environ.get("TDFO_TUNE_LIB"):          # a variant built by tools/build_variant.py
    _lib.LIB_PATH = Path(os.environ["TDFO_TUNE_LIB"]).resolve()

F, rows, D, P, B = (int(os.environ.get(k, d)) for k, d in (('F', 26), ('ROWS', 1_000_000), ('D', 128), ('P', 20), ('B', 16384)))
dev = torch.device("cuda")
gen = torch.Generator(device=dev).manual_seed(42)
w = torch.empty((F * rows, D), device=dev).uniform_(-0.001, 0.001)
grp = ops.TableGroup(weights=w, feat_row_base=torch.arange(F, device=dev) * rows,
                     feat_rows=torch.full((F,), rows, device=dev, dtype=torch.int64),
                     feat_out_col=torch.arange(F, device=dev) * D, out_stride=F * D,
                     state0=torch.zeros(F * rows, device=dev))
lengths = torch.full((F * B,), P, dtype=torch.int32, device=dev)
offsets = ops.complete_cumsum(lengths)
L = F * B * P
batches = [torch.randint(0, rows, (L,), device=dev, generator=gen) for _ in range(3)]
grad = torch.randn((B, F * D), device=dev, generator=gen)
opt = ops.make_optimizer_args(_lib.OPT_ROWWISE_ADAGRAD, 0.01, 1e-8)
lay = ops.backward_layout(L, F * rows)
out = torch.empty((B, F * D), device=dev)
worklist = torch.empty(ops.hot_row_worklist_entries(L), dtype=torch.int32, device=dev)
U = int(torch.unique(batches[0] + torch.arange(F, device=dev).repeat_interleave(B * P) * rows).numel())
upd_bytes = F * B * D * 4 + U * (2 * D * 4 + 8)
fwd_bytes = L * (D * 4 + 8) + (F * B + 1) * 8 + F * B * D * 4

def run(tag, n=10):
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(n)]
    for it in range(n + 2):
        e = ev[max(it - 2, 0)]
        idx = batches[it % 3]
        e[0].record(); ops.embedding_forward(grp, idx, offsets, B, 0, out=out)
        e[1].record(); k, v = ops.linearize_indices(grp, idx, offsets, B, 0)
        e[2].record(); sk, sv = ops.radix_sort_pairs(k, v, lay.key_bits, grp.err_flag)
        e[3].record(); uk, ss, nu, fv = ops.unique_segments(sk, grp.err_flag, sv)
        e[4].record(); ops.sparse_update(grp, grad, sv, uk, ss, fv, nu, offsets, B, L, 0, opt, worklist=worklist)
        e[5].record()
    torch.cuda.synchronize()
    ms = [statistics.mean(ev[i][j].elapsed_time(ev[i][j + 1]) for i in range(n)) for j in range(5)]
    print(f"{tag:28s} fwd {ms[0]:.3f} ({fwd_bytes/ms[0]/1e6:.0f} GB/s) lin {ms[1]:.3f} sort {ms[2]:.3f} uniq {ms[3]:.3f} "
          f"upd {ms[4]:.3f} ({upd_bytes/ms[4]/1e6:.0f} GB/s) total {sum(ms):.3f}", flush=True)

var = sys.argv[1] if len(sys.argv) > 1 else "TDFO_UPD_VARIANT"
for v in (sys.argv[2:] or ["0"]):
    os.environ[var] = v
    run(f"{var}={v}")
assert int(grp.err_flag.item()) == 0
