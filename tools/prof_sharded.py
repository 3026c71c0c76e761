"""Tuning helper (not product): torch.profiler breakdown of the sharded step on N GPUs (rank 0 prints)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import tdfo_b200 as t
from tdfo_b200 import distributed as td

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
F, rows, D, P, B = 26, 1_000_000, 128, 20, 16384
tables = [t.EmbeddingBagConfig(num_embeddings=rows, embedding_dim=D, name=f"t{i}", feature_names=[f"f{i}"]) for i in range(F)]
n_tw = F if os.environ.get("PLAN", "mixed") == "tw" else (F // world) * world
plan = {f"t{i}": (td.table_wise(i % world) if i < n_tw else td.row_wise()) for i in range(F)}
if os.environ.get("PLAN") == "cw":
    plan = td.balanced_plan(tables, world)
m = td.ShardedEmbeddingBagCollection(tables, plan, dev, init_on_device=True)
m.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=0.01)
m.train()
keys = [f"f{i}" for i in range(F)]
gen = torch.Generator(device=dev).manual_seed(rank)
lengths = torch.full((F * B,), P, dtype=torch.int32, device=dev)
vals = [torch.randint(0, rows, (F * B * P,), device=dev, generator=gen) for _ in range(3)]
grad = torch.randn((B, F * D), device=dev, generator=gen)
def step(i):
    out = m(t.KeyedJaggedTensor(keys, vals[i % 3], lengths=lengths, stride=B)).values()
    out.backward(grad)
for i in range(5): step(i)
torch.cuda.synchronize(); dist.barrier()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for i in range(5): step(i)
    torch.cuda.synchronize()
if rank == 0:
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=32, max_name_column_width=70), flush=True)
    # timeline of one step on rank 0: kernel name, start offset, duration (us)
    evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA),
                 key=lambda e: e.time_range.start)
    t0 = evs[0].time_range.start
    n_per = len(evs) // 5
    for e in evs[-n_per:]:
        print(f"{(e.time_range.start - t0):9.0f} {e.time_range.elapsed_us():8.0f}  {e.name[:90]}", flush=True)
os._exit(0)
