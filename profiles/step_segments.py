"""Where a multi-GPU step's time goes, measured with CUDA events on every rank (no nsys on this image):

    graph A  (rollout, value pass, scan, both backward passes)  |  all-reduce of [gradient ‖ previous KL]  |  graph B

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        profiles/step_segments.py [--steps 200]

The all-reduce segment of a rank = the collective's own latency + the time this rank waited for the slowest one; the
rank that arrives last sees (almost) only the latency, so  min over ranks  ~ latency  and  max - min  ~ skew between
ranks.  Rank 0 prints one JSON object.  BASELINE.json configs[2] per GPU (Pendulum, 65,536 envs x 20 steps)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from actor_critic_algs_on_tensorflow_b200 import trainer as TR  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=200)
ap.add_argument("--warmup", type=int, default=10)
ap.add_argument("--exchange", choices=["auto", "p2p", "nccl"], default="nccl",
                help="nccl: three segments; p2p: the exchange is a kernel inside the step's one graph, only step_ms is measured")
args = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))

marks = []
orig = TR.sync_exchange


def timed_exchange(buf, pg=None):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    orig(buf, pg)
    b.record()
    marks.append((a, b))
    return buf


TR.sync_exchange = timed_exchange
tr = TR.A2CTrainer("Pendulum-v0", n_envs=65536, horizon=20, seed=12321, rank=rank, world=world, mode="sync",
                   exchange=args.exchange)
for _ in range(args.warmup):
    tr.iteration()
torch.cuda.synchronize()
dist.barrier()
marks.clear()
steps = []
for _ in range(args.steps):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    tr.iteration()
    e.record()
    steps.append((s, e))
torch.cuda.synchronize()
seg = torch.zeros(4, dtype=torch.float64)
if marks:
    for (s, e), (a, b) in zip(steps, marks):
        seg += torch.tensor([s.elapsed_time(a), a.elapsed_time(b), b.elapsed_time(e), s.elapsed_time(e)],
                            dtype=torch.float64)
else:
    for s, e in steps:
        seg[3] += s.elapsed_time(e)
seg /= len(steps)
allseg = [torch.zeros(4, dtype=torch.float64, device="cuda") for _ in range(world)]
dist.all_gather(allseg, seg.cuda())
if rank == 0:
    m = torch.stack(allseg).cpu()
    names = ["graph_a_ms", "all_reduce_ms", "graph_b_ms", "step_ms"]
    out = {"n_gpus": world, "steps": args.steps, "exchange": tr.exchange,
           "per_rank": [{n: round(float(m[r, i]), 4) for i, n in enumerate(names)} for r in range(world)],
           "mean": {n: round(float(m[:, i].mean()), 4) for i, n in enumerate(names)},
           "min": {n: round(float(m[:, i].min()), 4) for i, n in enumerate(names)},
           "max": {n: round(float(m[:, i].max()), 4) for i, n in enumerate(names)}}
    out["collective_latency_ms"] = out["min"]["all_reduce_ms"]
    out["rank_skew_ms"] = round(out["max"]["all_reduce_ms"] - out["min"]["all_reduce_ms"], 4)
    print(json.dumps(out), flush=True)
dist.barrier()
dist.destroy_process_group()
