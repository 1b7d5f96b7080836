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


# ---------------------------------------------------------------------------------------------
# Limb-sharded key-switch (third mode): the one place where the path has a real exchange step
# ---------------------------------------------------------------------------------------------
def shard_layout(level, k, world, rank):
    """Pure host logic (no GPU): which Q / P limbs rank `rank` owns and the padded counts of the gathered buffers.
    Must agree with lsb::ShardPlan (csrc/lsb_shard.cu) -- checked against lsb_ks_shard_layout in the GPU tests."""
    l = level + 1
    own_q = list(range(rank, l, world))
    own_p = list(range(rank, k, world))
    return {"own_q": own_q, "own_p": own_p, "cq": -(-l // world), "cp": -(-k // world)}


def shard_poly(arr, own_q):
    """[..., l, N] -> [..., nq_own, N]: the owned Q limbs of a polynomial / ciphertext"""
    return arr[..., own_q, :].copy()


def shard_evk(evk, key_level, k, own_q, own_p):
    """[beta][2][key_level+1+k][N] -> [beta][2][nq_own+np_own][N] (owned Q limbs first, then owned P limbs)"""
    idx = list(own_q) + [key_level + 1 + j for j in own_p]
    return evk[:, :, idx, :].copy()


def unshard_ct(parts, level, world):
    """inverse of shard_poly over all ranks: list of [2][nq_own_r][N] -> [2][l][N]"""
    import numpy as np
    l = level + 1
    out = np.zeros((2, l, parts[0].shape[-1]), dtype=parts[0].dtype)
    for r, part in enumerate(parts):
        for u, i in enumerate(range(r, l, world)):
            out[:, i, :] = part[:, u, :]
    return out


class ShardedKeySwitch:
    """One rank of the limb-sharded hybrid key-switch: three local phases (C ABI lsb_ks_shard_*) around two
    all-gathers.  `gather(local, out)` fills `out` ([world, *local.shape]) with every rank's `local`; the default is
    torch.distributed.all_gather_into_tensor (NCCL over NVLink on the GPU box), tests pass an emulation."""

    def __init__(self, ctx, level, world, rank, device=None, gather=None):
        import ctypes as C
        import lattisense_b200 as L
        self.L, self.C, self.ctx = L, C, ctx
        self.level, self.world, self.rank = level, world, rank
        lay = shard_layout(level, ctx.k, world, rank)
        self.own_q, self.own_p, self.cq, self.cp = lay["own_q"], lay["own_p"], lay["cq"], lay["cp"]
        vals = [C.c_int() for _ in range(4)]
        oq = (C.c_int * max(1, len(self.own_q)))()
        op = (C.c_int * max(1, len(self.own_p)))()
        L._check(L.lib().lsb_ks_shard_layout(ctx.h, level, world, rank, *[C.byref(v) for v in vals], oq, op))
        got = ([v.value for v in vals], list(oq)[:len(self.own_q)], list(op)[:len(self.own_p)])
        want = ([len(self.own_q), len(self.own_p), self.cq, self.cp], self.own_q, self.own_p)
        if got != want:
            raise L.LsbError("shard layout mismatch between host and library: %r vs %r" % (got, want))
        self.nq_own, self.np_own = len(self.own_q), len(self.own_p)
        self.n_own = self.nq_own + self.np_own
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.gather = gather or self._nccl_gather
        N = ctx.N
        z = lambda *shape: torch.zeros(shape, dtype=torch.int64, device=self.device)
        self.coeff_own = z(self.cq, N)
        self.coeff_all = z(world, self.cq, N)
        self.acc_own = z(2, max(self.n_own, 1), N)
        self.accP_own = z(2, self.cp, N)
        self.accP_all = z(world, 2, self.cp, N)

    @staticmethod
    def _nccl_gather(local, out):
        # concatenation form ([world * rows, ...]): accepted by NCCL and gloo alike
        dist.all_gather_into_tensor(out.view(-1, *local.shape[1:]), local)

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def begin(self, c_own_ntt):
        L = self.L
        L._check(L.lib().lsb_ks_shard_begin(self.ctx.h, self.level, self.world, self.rank, c_own_ntt.data_ptr(),
                                            self.coeff_own.data_ptr(), 1, self._stream()))

    def inner(self, c_own_ntt, evk_own):
        L = self.L
        L._check(L.lib().lsb_ks_shard_inner(self.ctx.h, self.level, self.world, self.rank, self.coeff_all.data_ptr(),
                                            c_own_ntt.data_ptr(), evk_own.data_ptr(), self.acc_own.data_ptr(),
                                            self.accP_own.data_ptr(), 1, self._stream()))

    def end(self, addend_own, out_own):
        L = self.L
        L._check(L.lib().lsb_ks_shard_end(self.ctx.h, self.level, self.world, self.rank, self.accP_all.data_ptr(),
                                          self.acc_own.data_ptr(), addend_own.data_ptr() if addend_own is not None else None,
                                          out_own.data_ptr(), 1, self._stream()))

    # ---- fused compute + collective: the INTT kernels store into every rank's gathered buffer over peer memory ----
    def enable_p2p(self, group=None):
        """Re-home coeff_all / accP_all in torch symmetric memory and exchange the peer mappings (NVLink P2P).
        After this, run_p2p() needs no NCCL collective: only two cross-rank barriers per key-switch."""
        import torch.distributed._symmetric_memory as symm_mem
        n_c, n_p = self.coeff_all.numel(), self.accP_all.numel()
        buf = symm_mem.empty((n_c + n_p,), dtype=torch.int64, device=self.device)
        buf.zero_()
        group = group or dist.group.WORLD
        self._symm = symm_mem.rendezvous(buf, group.group_name)
        self._symm_buf = buf
        self.coeff_all = buf[:n_c].view(self.coeff_all.shape)
        self.accP_all = buf[n_c:].view(self.accP_all.shape)
        self.set_peers([int(p) for p in self._symm.buffer_ptrs], [int(p) + 8 * n_c for p in self._symm.buffer_ptrs])
        self._barrier = lambda ch: self._symm.barrier(channel=ch)

    def set_peers(self, coeff_ptrs, accp_ptrs, barrier=None):
        """peer tables: device pointers of every rank's coeff_all / accP_all as mapped in this process"""
        C = self.C
        self._peer_c = (C.c_void_p * self.world)(*coeff_ptrs)
        self._peer_p = (C.c_void_p * self.world)(*accp_ptrs)
        if barrier is not None:
            self._barrier = barrier

    def begin_p2p(self, c_own_ntt):
        L = self.L
        L._check(L.lib().lsb_ks_shard_begin_p2p(self.ctx.h, self.level, self.world, self.rank, c_own_ntt.data_ptr(),
                                                self._peer_c, self._stream()))

    def inner_p2p(self, c_own_ntt, evk_own):
        L = self.L
        L._check(L.lib().lsb_ks_shard_inner_p2p(self.ctx.h, self.level, self.world, self.rank, self.coeff_all.data_ptr(),
                                                c_own_ntt.data_ptr(), evk_own.data_ptr(), self.acc_own.data_ptr(),
                                                self._peer_p, self._stream()))

    def run_p2p(self, c_own_ntt, evk_own, addend_own=None):
        """same result as run(); barrier 0 also keeps a fast rank's next phase-2 stores out of accP_all while a slow
        rank still reads it in phase 3, barrier 1 does the same for coeff_all"""
        out = torch.empty((2, max(self.nq_own, 1), self.ctx.N), dtype=torch.int64, device=self.device)
        self.begin_p2p(c_own_ntt)
        self._barrier(0)
        self.inner_p2p(c_own_ntt, evk_own)
        self._barrier(1)
        self.end(addend_own, out)
        return out[:, :self.nq_own]

    def run(self, c_own_ntt, evk_own, addend_own=None):
        """c_own_ntt [nq_own][N], evk_own [beta][2][n_own][N], addend_own [2][nq_own][N] (int64 views of u64 residues,
        on this rank's GPU) -> out_own [2][nq_own][N]"""
        out = torch.empty((2, max(self.nq_own, 1), self.ctx.N), dtype=torch.int64, device=self.device)
        self.begin(c_own_ntt)
        self.gather(self.coeff_own, self.coeff_all)
        self.inner(c_own_ntt, evk_own)
        self.gather(self.accP_own, self.accP_all)
        self.end(addend_own, out)
        return out[:, :self.nq_own]


class NcclShardedKeySwitch:
    """Binding of the C++ driver of the limb-sharded key-switch (lsb_ks_shard_create / _run, csrc/lsb_shard_nccl.cu): the
    library owns the NCCL communicator, the exchange buffers and the CUDA graph; this class only distributes the NCCL
    unique id over torch.distributed (rendezvous plumbing) and passes device pointers."""

    def __init__(self, ctx, level, world, rank, max_batch=1, device=None):
        import ctypes as C
        import lattisense_b200 as L
        self.L, self.C, self.ctx = L, C, ctx
        self.level, self.world, self.rank, self.max_batch = level, world, rank, max_batch
        lay = shard_layout(level, ctx.k, world, rank)
        self.own_q, self.own_p = lay["own_q"], lay["own_p"]
        self.nq_own, self.np_own = len(self.own_q), len(self.own_p)
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        idbuf = (C.c_uint8 * 128)()
        if rank == 0:
            L._check(L.lib().lsb_nccl_unique_id(idbuf))
        if world > 1:
            t = torch.tensor(list(idbuf), dtype=torch.uint8, device=self.device)
            dist.broadcast(t, src=0)
            idbuf = (C.c_uint8 * 128)(*t.cpu().tolist())
        self.h = L.lib().lsb_ks_shard_create(ctx.h, level, world, rank, idbuf, max_batch)
        if not self.h:
            raise L.LsbError(L.lib().lsb_ks_shard_last_error().decode())

    def run(self, c_own_ntt, evk_own, addend_own=None, out=None, stream=None):
        """c_own_ntt [B][nq_own][N], evk_own [beta][2][n_own][N], addend_own [B][2][nq_own][N] -> out [B][2][nq_own][N]"""
        L = self.L
        B = c_own_ntt.shape[0]
        if out is None:
            out = torch.empty((B, 2, max(self.nq_own, 1), self.ctx.N), dtype=torch.int64, device=self.device)
        s = stream if stream is not None else torch.cuda.current_stream(self.device).cuda_stream
        rc = L.lib().lsb_ks_shard_run(self.h, c_own_ntt.data_ptr(), evk_own.data_ptr(), addend_own.data_ptr() if addend_own is not None else None,
                                      out.data_ptr(), B, s)
        if rc != 0:
            raise L.LsbError(L.lib().lsb_ks_shard_last_error().decode())
        return out

    def all_gather_bytes(self):
        return int(self.L.lib().lsb_ks_shard_all_gather_bytes(self.h))

    def close(self):
        if getattr(self, "h", None):
            self.L.lib().lsb_ks_shard_destroy(self.h)
            self.h = None
