"""torchrun smoke: every rank reports its environment and does one NCCL all-reduce (used to debug launch problems)"""
import os
import sys

print("diag: pid", os.getpid(), "RANK", os.environ.get("RANK"), "LOCAL_RANK", os.environ.get("LOCAL_RANK"), "WORLD_SIZE",
      os.environ.get("WORLD_SIZE"), "argv", sys.argv, flush=True)
import torch
import torch.distributed as dist

lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
t = torch.ones(4, device="cuda") * (dist.get_rank() + 1)
dist.all_reduce(t)
print("diag: rank", dist.get_rank(), "allreduce", t.tolist(), flush=True)
dist.destroy_process_group()
