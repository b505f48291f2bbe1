"""Sweep of the Ulysses pipelining knobs on the Wan layer (development tool; run under torchrun on the GPU box):
    python -m torch.distributed.run --nproc-per-node P --master-addr 127.0.0.1 tools/sp_sweep.py "1:0,2:0,2:8,2:16"
Each item is head_chunks:comm_sms.  Prints one JSON line per setting (bench.run_sp's record)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402

world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)


def barrier():
    dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


for item in sys.argv[1].split(","):
    c, s = (int(x) for x in item.split(":"))
    transport = sys.argv[2] if len(sys.argv) > 2 else "nccl"
    rec = bench.run_sp(dev, world, rank, barrier, max_over_ranks, transport=transport, chunks=c, comm_sms=s, reps=5)
    if rank == 0:
        print(json.dumps({"chunks": c, "comm_sms": s, "ms": rec["ms"], "comm_ms": rec["comm_ms"],
                          "one_gpu_ms": rec["one_gpu_ms"], "efficiency": rec["efficiency"], "tops": rec["tops"]}), flush=True)
dist.barrier()
dist.destroy_process_group()
