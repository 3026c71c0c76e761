"""Measurement helper (not product): does the id-only half of the backward (linearise, sort, run-length) hide behind
the pooled lookup when it runs on a second stream?  BASELINE config 2, device-resident, three schedules:
  serial   lookup (emits the pairs) -> sort/unique -> update          (the product path)
  side-1st prepare on a side stream enqueued BEFORE the lookup, update waits for both
  side-2nd prepare on a side stream enqueued AFTER the lookup"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tdfo_b200 import ops, _lib

F, rows, D, P, B = 26, 1_000_000, 128, 20, 16384
dev = torch.device("cuda")
gen = torch.Generator(device=dev).manual_seed(42)
w = torch.empty((F * rows, D), device=dev).uniform_(-0.001, 0.001)
grp = ops.TableGroup(weights=w, feat_row_base=torch.arange(F, device=dev) * rows,
                     feat_rows=torch.full((F,), rows, device=dev, dtype=torch.int64),
                     feat_out_col=torch.arange(F, device=dev) * D, out_stride=F * D,
                     state0=torch.zeros(F * rows, device=dev))
offsets = ops.complete_cumsum(torch.full((F * B,), P, dtype=torch.int32, device=dev))
L = F * B * P
batches = [torch.randint(0, rows, (L,), device=dev, generator=gen) for _ in range(4)]
grad = torch.randn((B, F * D), device=dev, generator=gen)
opt = ops.make_optimizer_args(_lib.OPT_ROWWISE_ADAGRAD, 0.01, 1e-8)
scratch = torch.empty(ops.backward_layout(L, F * rows).total_bytes, dtype=torch.uint8, device=dev)
out = torch.empty((B, F * D), device=dev)
side = torch.cuda.Stream(device=dev)
main = torch.cuda.current_stream(dev)


def serial(i):
    idx = batches[i % 4]
    ops.embedding_forward(grp, idx, offsets, B, 0, out=out, backward_scratch=scratch)
    ops.embedding_backward_update(grp, grad, idx, offsets, B, 0, opt, scratch=scratch, keys_in_scratch=True)


def overlapped(i, side_first):
    idx = batches[i % 4]
    side.wait_stream(main)                      # the previous update has released the scratch
    if side_first:
        with torch.cuda.stream(side):
            ops.embedding_backward_prepare(grp, idx, offsets, B, 0, scratch)
    ops.embedding_forward(grp, idx, offsets, B, 0, out=out)
    if not side_first:
        with torch.cuda.stream(side):
            ops.embedding_backward_prepare(grp, idx, offsets, B, 0, scratch)
    main.wait_stream(side)
    ops.embedding_backward_update(grp, grad, idx, offsets, B, 0, opt, scratch=scratch, keys_in_scratch="prepared")


def timeit(name, fn, n=20):
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(5 + i)
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:10s} {e0.elapsed_time(e1) / n:.3f} ms/step", flush=True)


for _ in range(2):
    timeit("serial", serial)
    timeit("side-1st", lambda i: overlapped(i, True))
    timeit("side-2nd", lambda i: overlapped(i, False))
assert int(grp.err_flag.item()) == 0
