"""Tuning helper (not product): torch.profiler breakdown of one sharded step of a BASELINE config on N GPUs
(rank 0 prints).  CONFIG=2|3|4|5 (default 2), PLAN=tw|mixed|cw overrides config 2's plan, PREFETCH=1."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import tdfo_b200 as t
import bench_workloads as bw
from tdfo_b200 import distributed as td

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
cfg = int(os.environ.get("CONFIG", "2"))
sp = bw.spec(cfg, world)
rows, D, B = sp["rows"], sp["D"], sp["B"]
F = len(rows)
tables = bw.make_tables(t, sp)
override = os.environ.get("PLAN") if cfg == 2 else None
plan = bw.make_plan(td, sp, tables, world, None if override in (None, "cw") else override)
batches = [bw.device_batch(torch, sp, rows, B, dev, seed=42 + rank + 1000 * bi) for bi in range(2)]
tot = torch.tensor([float(max(b[1].numel() for b in batches))], dtype=torch.float64, device=dev)
dist.all_reduce(tot)
m = td.ShardedEmbeddingBagCollection(tables, plan, dev, init_on_device=True, max_ids_per_rank=int(tot.item()) + 1024)
m.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=0.01)
m.train()
keys = [f"f{i}" for i in range(F)]
grad = torch.randn((B, F * D), device=dev)
prefetch = os.environ.get("PREFETCH") == "1"
nxt = [None]
def kjt(i):
    return t.KeyedJaggedTensor(keys, batches[i % 2][1], lengths=batches[i % 2][0], stride=B)
def step(i, last):
    k = nxt[0] if nxt[0] is not None else kjt(i)
    nxt[0] = None
    out = m(k).values()
    if prefetch and not last:
        nxt[0] = kjt(i + 1)
        m.prefetch_input_dist(nxt[0])
    out.backward(grad)
for i in range(4): step(i, i == 3)
torch.cuda.synchronize(); dist.barrier()
from torch.profiler import profile, ProfilerActivity
N = 4
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for i in range(N): step(i, i == N - 1)
    torch.cuda.synchronize()
if rank == 0:
    print(f"# {sp['name']}; world {world}; prefetch {prefetch}")
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=32, max_name_column_width=70), flush=True)
    # timeline of one step on rank 0: kernel name, start offset, duration (us)
    evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA),
                 key=lambda e: e.time_range.start)
    t0 = evs[0].time_range.start
    n_per = len(evs) // N
    for e in evs[-2 * n_per:-n_per]:
        print(f"{(e.time_range.start - t0):9.0f} {e.time_range.elapsed_us():8.0f}  {e.name[:90]}", flush=True)
os._exit(0)
