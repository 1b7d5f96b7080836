"""One-process-per-GPU plumbing for the batch-sharded path: independent ciphertext batches /
DAG components are sharded across ranks with NO data-path collective; torch.distributed is used only for the
rendezvous, the barriers around the timed region and the max-over-ranks of the device time.

Works with backend "nccl" (GPU box, under torchrun) and "gloo" (CPU tests, world_size 2).
"""
import os

import torch
import torch.distributed as dist


class Ranks:
    def __init__(self, backend=None):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        self.device = torch.device("cuda", self.local) if self.backend == "nccl" else torch.device("cpu")
        if self.world > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            if self.backend == "nccl":
                torch.cuda.set_device(self.local)
                dist.init_process_group("nccl", device_id=self.device)
            else:
                dist.init_process_group("gloo")

    def barrier(self):
        if self.backend == "nccl":
            torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
            if self.backend == "nccl":
                torch.cuda.synchronize()

    def max(self, value):
        """max over ranks of a python float (the device time of the slowest rank)"""
        t = torch.tensor([float(value)], dtype=torch.float64, device=self.device)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum(self, value):
        t = torch.tensor([float(value)], dtype=torch.float64, device=self.device)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def shard(self, n_units):
        """contiguous share of n_units independent units owned by this rank (strong-scaling helper)"""
        base, rem = divmod(n_units, self.world)
        lo = self.rank * base + min(self.rank, rem)
        return lo, lo + base + (1 if self.rank < rem else 0)

    def whole_job_rate(self, units_this_rank, seconds_this_rank):
        """units all ranks processed / slowest rank's time"""
        total = self.sum(units_this_rank)
        return total / self.max(seconds_this_rank)

    def close(self):
        if self.world > 1 and dist.is_initialized():
            dist.destroy_process_group()
