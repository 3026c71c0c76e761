"""Tuning helper (not product): times tdfo_radix_sort_pairs alone on config-2-sized keys (8.5M pairs, 25 key bits).
TDFO_TUNE_LIB selects a variant built by tools/build_variant.py (timing-only variants give wrong orders)."""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pathlib import Path
from tdfo_b200 import ops, _lib
if os.environ.get("TDFO_TUNE_LIB"):
    _lib.LIB_PATH = Path(os.environ["TDFO_TUNE_LIB"]).resolve()
dev = torch.device("cuda")
n, bits = int(os.environ.get("N", 8519680)), int(os.environ.get("BITS", 25))
g = torch.Generator(device=dev).manual_seed(1)
keys0 = torch.randint(0, 26_000_000 if bits == 25 else (1 << bits), (n,), device=dev, generator=g, dtype=torch.int64).to(torch.int32)
vals0 = torch.arange(n, device=dev, dtype=torch.int32)
flag = torch.zeros(1, dtype=torch.int32, device=dev)
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(12)]
for a, b in ev:
    k, v = keys0.clone(), vals0.clone()
    a.record(); ko, vo = ops.radix_sort_pairs(k, v, bits, flag); b.record()
torch.cuda.synchronize()
ms = statistics.mean(a.elapsed_time(b) for a, b in ev[2:])
ok = bool((ko[1:].long() >= ko[:-1].long()).all())
print(f"{sys.argv[1] if len(sys.argv) > 1 else 'base':14s} n={n} bits={bits} sort {ms:.3f} ms sorted={ok}", flush=True)
