#!/usr/bin/env python
"""bench.py — BASELINE.json metric on its quoted configuration.

Workload (BASELINE.json configs[1]): CKKS mult_relin + rescale, n = 16384, level 3 -> 2
(parameter set frontend/parameter.json:95-114), a batch of 1024 ciphertexts per GPU.  One "step" =
one pass of the hot path over the batch.  `value` = ciphertext ops/s with inputs resident in HBM;
`e2e` = the same through the C ABI with HOST buffers (pinned H2D of both input batches and D2H of
the result inside the timed region).  Inputs (2 GiB per step) are larger than L2, so no L2 flush
is needed between iterations.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]

Under torchrun (N > 1) every rank drives its own GPU with its own batch (weak scaling, no data-path
collective: ciphertext batches are independent); time = max over ranks.
`--impl reference` times the CPU path (oracle restatement of the reference's Lattigo path; the
reference's own CPU library cannot be built here: Go + the lattigo submodule are absent) on the
host cores for the same metric.
"""
import argparse
import json
import multiprocessing as mp
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

METRIC = "ckks_mult_relin_rescale_ops_per_s"
UNIT = "ciphertext ops/s"
LEVEL = 3


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region"""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                self.rows.append([x.strip() for x in out.stdout.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        sm = sorted(int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU arm (oracle = restatement of the reference's Lattigo CPU path)
# ---------------------------------------------------------------------------------------------
_W = {}


def _cpu_init(variant="optimised"):
    import fhe_ref as R
    o = R.Oracle(R.CKKS_N14)
    _W["variant"] = variant
    if variant == "optimised":
        _W["f"] = R.FastOracle(R.CKKS_N14)
    rng = np.random.default_rng(os.getpid())
    q, p = R.CKKS_N14.q, R.CKKS_N14.p
    l = LEVEL + 1
    mk = lambda mods, lead: np.stack([rng.integers(0, m, size=lead + (o.N,), dtype=np.uint64) for m in mods], axis=-2)
    _W["o"] = o
    _W["a"], _W["b"] = mk(q[:l], (2,)), mk(q[:l], (2,))
    _W["rlk"] = np.ascontiguousarray(mk(q[:l] + p, (o.p.beta(LEVEL), 2)))


def _cpu_work(n):
    o = _W["o"]
    if _W["variant"] == "optimised":
        return _W["f"].time_mult_relin_rescale(_W["a"], _W["b"], _W["rlk"], n)
    t0 = time.perf_counter()
    for _ in range(n):
        o.rescale(o.relin(o.mult(_W["a"], _W["b"]), _W["rlk"]))
    return time.perf_counter() - t0


def cpu_ops_per_s(cores, ops_per_core, variant="optimised"):
    """all cores, one CPU context per process (the reference CPU runner uses min(32, hw) pool threads,
    mega_ag_runners/cpu/cpu_wrapper.cpp:57).  variant "optimised": oracle/orc_fast.hpp (Shoup/Barrett lazy arithmetic,
    -O3; limb-equal to the simple oracle, tests/test_oracle_cpu.py) - the figure the ratios are quoted against;
    "oracle": the simple division-based restatement the parity tests use."""
    import fhe_ref as R
    R.build_oracle()
    cores = min(cores, 32)
    with mp.get_context("fork").Pool(cores, initializer=_cpu_init, initargs=(variant,)) as pool:
        pool.map(_cpu_work, [1] * cores)  # warm-up
        t0 = time.perf_counter()
        pool.map(_cpu_work, [ops_per_core] * cores)
        dt = time.perf_counter() - t0
    return cores * ops_per_core / dt, dt


def run_reference(args):
    """--impl reference: the CPU path on the host cores, same metric/config; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = min(len(os.sched_getaffinity(0)), 32)
    per_core = 100
    steps_val = []
    for _ in range(min(args.warmup, 2)):
        cpu_ops_per_s(cores, 4)
    t_all = 0.0
    for _ in range(args.steps):
        v, dt = cpu_ops_per_s(cores, per_core)
        steps_val.append(v)
        t_all += dt
    value = float(np.mean(steps_val))
    simple, _ = cpu_ops_per_s(cores, 10, "oracle")
    sample = "%d mult_relin_rescale ops per step (%d per core) of the %d-ciphertext batch" % (
        cores * per_core, per_core, args.batch)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_all / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "CKKS mult_relin+rescale n=16384 level 3->2 (BASELINE configs[1])",
                   "batch_per_gpu": args.batch, "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "variant": "optimised", "sample": sample,
                         "oracle_variant_value": simple},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU baseline is a port of the Lattigo path (Go toolchain + lattigo submodule absent): variant 'optimised' = "
                "oracle/orc_fast.hpp (Shoup/Barrett lazy NTT + MAC, -O3, limb-equal to the oracle); 'oracle_variant_value' = the "
                "simple division-based oracle the parity tests use",
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# API-level arm: the same workload through the reference-facing boundary (run_fhe_gpu_task)
# ---------------------------------------------------------------------------------------------
def write_cfg2_task(path, B, P):
    """mega_ag.json / task_signature.json of B independent z_i = rescale(mult_relin(x_i, y_i)) at level 3 in the
    format frontend/custom_task.py emits (field for field the committed fixture tests/golden/graphs/cfg2_ckks_mult_x8,
    which the reference frontend generated; here only the copy count differs)."""
    os.makedirs(path, exist_ok=True)
    ct = lambda i, lv, deg=1, t="ct": {"id": i, "type": t, "level": lv, "degree": deg, "is_ntt": True, "is_mform": False,
                                       "poly1_rns_sp_decomped": False}
    data, comp = {}, {}
    nd = 0
    xs, ys, zs = [], [], []
    for i in range(B):
        data[str(nd)] = ct("x_%d" % i, LEVEL); xs.append(nd); nd += 1
    for i in range(B):
        data[str(nd)] = ct("y_%d" % i, LEVEL); ys.append(nd); nd += 1
    rlk = nd
    data[str(nd)] = {"id": "rlk_ntt", "type": "rlk", "level": LEVEL, "degree": 1, "is_ntt": True, "is_mform": True, "sp_level": len(P.p) - 1}
    nd += 1
    nc = 0
    for i in range(B):
        c3, r = nd, nd + 1
        z = nd + 2
        nd += 3
        data[str(c3)] = ct("t3_%d" % i, LEVEL, 2, "ct3")
        data[str(r)] = ct("t2_%d" % i, LEVEL)
        data[str(z)] = ct("z_%d" % i, LEVEL - 1)
        comp[str(nc)] = {"id": "m%d" % i, "type": "mult", "inputs": [xs[i], ys[i]], "outputs": [c3]}
        comp[str(nc + 1)] = {"id": "r%d" % i, "type": "relin", "inputs": [c3, rlk], "outputs": [r]}
        comp[str(nc + 2)] = {"id": "s%d" % i, "type": "rescale", "inputs": [r], "outputs": [z]}
        nc += 3
        zs.append(z)
    g = {"name": "Acc task", "algorithm": "CKKS", "data": data, "compute": comp, "inputs": xs + ys + [rlk], "outputs": zs,
         "offline_inputs": [], "parameter": {"n": P.N, "max_level": len(P.q) - 1, "q": P.q, "p": P.p, "slots": P.N // 2, "scale": P.scale}}
    sig = {"algorithm": "CKKS", "key": {"rlk": LEVEL, "glk": {}},
           "online": [{"id": "in_x_list", "type": "ct", "size": [B], "level": LEVEL, "phase": "in"},
                      {"id": "in_y_list", "type": "ct", "size": [B], "level": LEVEL, "phase": "in"},
                      {"id": "out_z_list", "type": "ct", "size": [B], "level": LEVEL - 1, "phase": "out"}], "offline": []}
    with open(os.path.join(path, "mega_ag.json"), "w") as f:
        json.dump(g, f)
    with open(os.path.join(path, "task_signature.json"), "w") as f:
        json.dump(sig, f)


def write_benchmark_gpu_task(path, kind, B, P, level=3):
    """The three graphs of the reference's own GPU benchmark (examples/benchmark_gpu/benchmark_gpu.py:26-75: 1024 x BFV
    mult_relin, 1024 x CKKS mult_relin, 1024 x BFV rotate_cols(x, 1), all n = 16384, level 3) in the format the frontend
    emits (compare tests/golden/graphs/cfg1_bfv_mult and bfv16k_rotate_col_l3, which the reference frontend generated)."""
    os.makedirs(path, exist_ok=True)
    bfv = kind.startswith("bfv")
    ct = lambda i, deg=1, t="ct": {"id": i, "type": t, "level": level, "degree": deg, "is_ntt": not bfv, "is_mform": False,
                                   "poly1_rns_sp_decomped": False}
    key = lambda i, t: {"id": i, "type": t, "level": level, "degree": 1, "is_ntt": True, "is_mform": True, "sp_level": len(P.p) - 1}
    data, comp, nd = {}, {}, 0
    xs, ys, zs = [], [], []
    for i in range(B):
        data[str(nd)] = ct("x_%d" % i); xs.append(nd); nd += 1
    if kind.endswith("mult_relin"):
        for i in range(B):
            data[str(nd)] = ct("y_%d" % i); ys.append(nd); nd += 1
        k = nd
        data[str(nd)] = key("rlk_ntt", "rlk"); nd += 1
        for i in range(B):
            data[str(nd)] = ct("t3_%d" % i, 2, "ct3")
            data[str(nd + 1)] = ct("z_%d" % i)
            comp[str(2 * i)] = {"id": "m%d" % i, "type": "mult", "inputs": [xs[i], ys[i]], "outputs": [nd]}
            comp[str(2 * i + 1)] = {"id": "r%d" % i, "type": "relin", "inputs": [nd, k], "outputs": [nd + 1]}
            zs.append(nd + 1)
            nd += 2
        keys = {"rlk": level, "glk": {}}
        online_in = [("xs", B), ("ys", B)]
    else:
        k = nd
        data[str(nd)] = dict(key("glk_ntt_col_5", "glk"), galois_element=5); nd += 1
        for i in range(B):
            data[str(nd)] = ct("y_%d" % i)
            comp[str(i)] = {"id": "c%d" % i, "type": "rotate_col", "inputs": [xs[i], k], "outputs": [nd], "step": 1}
            zs.append(nd)
            nd += 1
        keys = {"rlk": -1, "glk": {"5": level}}
        online_in = [("xs", B)]
    par = {"n": P.N, "max_level": len(P.q) - 1, "q": P.q, "p": P.p}
    par.update({"t": P.t} if bfv else {"slots": P.N // 2, "scale": P.scale})
    g = {"name": "Acc task", "algorithm": "BFV" if bfv else "CKKS", "data": data, "compute": comp, "inputs": xs + ys + [k],
         "outputs": zs, "offline_inputs": [], "parameter": par}
    sig = {"algorithm": g["algorithm"], "key": keys,
           "online": [{"id": n, "type": "ct", "size": [c], "level": level, "phase": "in"} for n, c in online_in] +
                     [{"id": "out", "type": "ct", "size": [B], "level": level, "phase": "out"}], "offline": []}
    with open(os.path.join(path, "mega_ag.json"), "w") as f:
        json.dump(g, f)
    with open(os.path.join(path, "task_signature.json"), "w") as f:
        json.dump(sig, f)


def reference_benchmark_gpu(device, B=1024, level=3):
    """examples/benchmark_gpu: wall time of one run_fhe_gpu_task call on a warm task handle (what benchmark_gpu.cpp:50 prints:
    EXPORT + H2D + kernels + D2H + IMPORT), host limb buffers in and out, random residues for ciphertexts and keys (throughput
    does not depend on the values; parity of these node types: tests/test_runner_gpu.py)."""
    import ctypes as C
    import shutil
    import tempfile
    import torch
    import fhe_ref as R
    import graph_eval as GE
    out = {}
    u64p = C.POINTER(C.c_uint64)
    ia = lambda v: (C.c_int * len(v))(*v)

    def limbs(lead, mods, n):
        t = torch.empty(lead + (len(mods), n), dtype=torch.int64)
        for i, m in enumerate(mods):
            t[..., i, :] = torch.randint(0, m, lead + (n,), dtype=torch.int64)
        return t

    for kind, P in (("bfv_mult_relin", R.BFV_N14), ("ckks_mult_relin", R.CKKS_N14), ("bfv_rotate_col", R.BFV_N14)):
        d = tempfile.mkdtemp(prefix="lsb_refbench_")
        try:
            write_benchmark_gpu_task(d, kind, B, P, level)
            l = level + 1

            def handle(ptr, k, deg=1):
                t = GE.THandle()
                t.kind, t.level, t.degree, t.n, t.k = k, level, deg, P.N, len(P.p)
                t.data = C.cast(ptr, u64p)
                return t

            ct_bytes = 2 * l * P.N * 8
            groups = [limbs((B, 2), P.q[:l], P.N)]
            if kind.endswith("mult_relin"):
                groups.append(limbs((B, 2), P.q[:l], P.N))
            evk = limbs((P.beta(level), 2), P.q[:l] + P.p, P.N)
            res = torch.empty((B, 2, l, P.N), dtype=torch.int64)
            h_in = [handle(g.data_ptr() + i * ct_bytes, 0) for g in groups for i in range(B)]
            keep = []
            if kind.endswith("mult_relin"):
                h_in.append(handle(evk.data_ptr(), 3))
                meta_in = (ia([GE.TYPE_CIPHERTEXT] * 2 + [GE.TYPE_RELIN_KEY]), ia([level] * 3), ia([B, B, 1]))
            else:
                sub = handle(evk.data_ptr(), 3)
                ga = (C.c_uint64 * 1)(5)
                parr = (C.POINTER(GE.THandle) * 1)(C.pointer(sub))
                hs = GE.THandle()
                hs.kind, hs.n, hs.k, hs.n_keys = 4, P.N, len(P.p), 1
                hs.galois_elements = C.cast(ga, u64p)
                hs.keys = parr
                keep += [sub, ga, parr]
                h_in.append(hs)
                meta_in = (ia([GE.TYPE_CIPHERTEXT, GE.TYPE_GALOIS_KEY]), ia([level, -1]), ia([B, 1]))
            h_out = [handle(res.data_ptr() + i * ct_bytes, 0) for i in range(B)]
            meta_out = (ia([GE.TYPE_CIPHERTEXT]), ia([level]), ia([B]))
            in_ptrs = (C.POINTER(GE.THandle) * len(h_in))(*[C.pointer(x) for x in h_in])
            out_ptrs = (C.POINTER(GE.THandle) * len(h_out))(*[C.pointer(x) for x in h_out])
            prog = (C.c_int * 3)()
            h = GE.harness()
            task = h.th_open(d.encode(), 1)
            if not task:
                raise RuntimeError("th_open: " + h.th_last_error().decode())
            try:
                def run():
                    rc = h.th_run(task, len(meta_in[0]), meta_in[0], meta_in[1], meta_in[2], in_ptrs, 1, meta_out[0], meta_out[1],
                                  meta_out[2], out_ptrs, device, prog)
                    if rc != 0:
                        raise RuntimeError("run_fhe_gpu_task failed: " + h.th_last_error().decode())
                res.fill_(-1)
                run()  # warm-up: context, tables, pools, key upload
                if bool((res == -1).any()):
                    raise RuntimeError("the run left output limbs unwritten")
                best = 1e30
                for _ in range(3):
                    t0 = time.perf_counter()
                    run()
                    best = min(best, time.perf_counter() - t0)
                out[kind + "_x%d_ms" % B] = 1e3 * best
                out[kind + "_ops_per_s"] = B / best
            finally:
                h.th_close(task)
            del groups, evk, res
        except Exception as ex:  # a side measurement never costs the headline line
            out[kind + "_error"] = str(ex)[:200]
        finally:
            shutil.rmtree(d, ignore_errors=True)
    out["note"] = ("the reference's examples/benchmark_gpu graphs (n=16384, level 3, %d ops each) through run_fhe_gpu_task, "
                   "best of 3 wall-clock runs on a warm handle, host buffers in and out" % B)
    return out


class RunnerArm:
    """Drives create_/bind_/run_/release_fhe_gpu_task the way FheTaskGpu does (cxx_sdk_v2/cxx_fhe_task_gpu.cpp:69-117),
    with the test plug-in's EXPORT/IMPORT executors over host limb buffers (tests/harness/harness.cpp)."""

    def __init__(self, P, B, a_host, b_host, rlk_host, out_host, device, tag):
        import ctypes as C
        import tempfile
        import graph_eval as GE
        self.C, self.GE, self.B, self.device = C, GE, B, device
        self.dir = tempfile.mkdtemp(prefix="lsb_cfg2_%s_" % tag)
        write_cfg2_task(self.dir, B, P)
        self.h = GE.harness()
        self.task = self.h.th_open(self.dir.encode(), 1)
        if not self.task:
            raise RuntimeError("th_open: " + self.h.th_last_error().decode())
        u64p = C.POINTER(C.c_uint64)

        def handle(ptr, kind, level, degree):
            t = GE.THandle()
            t.kind, t.level, t.degree, t.n, t.k = kind, level, degree, P.N, len(P.p)
            t.data = C.cast(ptr, u64p)
            return t

        l = LEVEL + 1
        ct_bytes = 2 * l * P.N * 8
        self.h_in = [handle(a_host.data_ptr() + i * ct_bytes, 0, LEVEL, 1) for i in range(B)]
        self.h_in += [handle(b_host.data_ptr() + i * ct_bytes, 0, LEVEL, 1) for i in range(B)]
        self.h_in.append(handle(rlk_host.data_ptr(), 3, LEVEL, 1))
        out_bytes = 2 * (l - 1) * P.N * 8
        self.h_out = [handle(out_host.data_ptr() + i * out_bytes, 0, LEVEL - 1, 1) for i in range(B)]
        ia = lambda v: (C.c_int * len(v))(*v)
        self.in_meta = (ia([GE.TYPE_CIPHERTEXT, GE.TYPE_CIPHERTEXT, GE.TYPE_RELIN_KEY]), ia([LEVEL, LEVEL, LEVEL]), ia([B, B, 1]))
        self.out_meta = (ia([GE.TYPE_CIPHERTEXT]), ia([LEVEL - 1]), ia([B]))
        self.in_ptrs = (C.POINTER(GE.THandle) * len(self.h_in))(*[C.pointer(x) for x in self.h_in])
        self.out_ptrs = (C.POINTER(GE.THandle) * len(self.h_out))(*[C.pointer(x) for x in self.h_out])
        self.prog = (C.c_int * 3)()

    def step(self):
        rc = self.h.th_run(self.task, 3, self.in_meta[0], self.in_meta[1], self.in_meta[2], self.in_ptrs, 1, self.out_meta[0],
                           self.out_meta[1], self.out_meta[2], self.out_ptrs, self.device, self.prog)
        if rc != 0:
            raise RuntimeError("run_fhe_gpu_task failed: " + self.h.th_last_error().decode())

    def stats(self):
        C = self.C
        st = [C.c_longlong() for _ in range(3)]
        self.GE.runner_lib().lsb_gpu_task_last_stats(self.task, *[C.byref(x) for x in st])
        return tuple(x.value for x in st)

    def host_profile(self):
        C = self.C
        L = self.GE.runner_lib()
        ms = (C.c_double * 6)()
        L.lsb_gpu_task_host_profile.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.lsb_gpu_task_host_profile(self.task, ms)
        tr = [C.c_longlong() for _ in range(4)]
        L.lsb_gpu_task_transfer_stats(self.task, *[C.byref(x) for x in tr])
        d = dict(zip(["wall_ms", "export_ms_sum", "stage_ms_sum", "ring_wait_ms_sum", "d2h_wait_ms_sum", "import_ms_sum"], [round(x, 2) for x in ms]))
        d.update(dict(zip(["h2d_bytes", "d2h_bytes", "zero_copy_bytes", "key_bytes_cached"], [x.value for x in tr])))
        return d

    def close(self):
        import shutil
        self.h.th_close(self.task)
        shutil.rmtree(self.dir, ignore_errors=True)


def run_ours(args):
    import ctypes as C
    import torch
    import fhe_ref as R
    import lattisense_b200 as L

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    P = R.CKKS_N14
    N, l, B = P.N, LEVEL + 1, args.batch
    ctx = L.Context(P.scheme, P.logN, P.q, P.p, 0, device=local)
    stream = torch.cuda.current_stream().cuda_stream
    ctx.stream = stream
    lib = L.lib()

    def rand_limbs(lead, mods):
        t = torch.empty(lead + (len(mods), N), dtype=torch.int64, device=dev)
        for i, m in enumerate(mods):
            t[..., i, :] = torch.randint(0, m, lead + (N,), dtype=torch.int64, device=dev)
        return t

    torch.manual_seed(1234 + rank)
    a_dev = rand_limbs((B, 2), P.q[:l])
    b_dev = rand_limbs((B, 2), P.q[:l])
    rlk_dev = rand_limbs((P.beta(LEVEL), 2), P.q[:l] + P.p)
    out_dev = torch.empty((B, 2, l - 1, N), dtype=torch.int64, device=dev)
    a_host = torch.empty(a_dev.shape, dtype=torch.int64).pin_memory()
    b_host = torch.empty(b_dev.shape, dtype=torch.int64).pin_memory()
    out_host = torch.empty(out_dev.shape, dtype=torch.int64).pin_memory()
    a_host.copy_(a_dev)
    b_host.copy_(b_dev)
    torch.cuda.synchronize()

    def step_dev():
        L._check(lib.lsb_mult_relin_rescale(ctx.h, LEVEL, a_dev.data_ptr(), b_dev.data_ptr(), rlk_dev.data_ptr(),
                                            LEVEL, out_dev.data_ptr(), B, stream))

    def step_e2e():
        L._check(lib.lsb_upload(ctx.h, a_dev.data_ptr(), a_host.data_ptr(), a_host.numel(), stream))
        L._check(lib.lsb_upload(ctx.h, b_dev.data_ptr(), b_host.data_ptr(), b_host.numel(), stream))
        step_dev()
        L._check(lib.lsb_download(ctx.h, out_host.data_ptr(), out_dev.data_ptr(), out_host.numel(), stream))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(args.warmup, 3)):
        step_dev()
    sampler = ClockSampler(local)
    sampler.start()
    n0 = ctx.launch_count()
    ms = timed(step_dev, args.steps)
    launches = ctx.launch_count() - n0
    clocks = sampler.summary()
    value = world * B * args.steps / (ms / 1e3)

    # end-to-end through the kernel-level C ABI with host buffers
    step_e2e()
    e2e_steps = max(1, min(args.steps, 5))
    ms_e2e = timed(step_e2e, e2e_steps)
    e2e_value = world * B * e2e_steps / (ms_e2e / 1e3)

    # end-to-end through the reference boundary: run_fhe_gpu_task on a 3*B-node mega_ag graph, host limb buffers in,
    # EXPORT -> pinned staging -> H2D -> batched kernels -> D2H -> IMPORT into host buffers out (wall clock: the call is
    # synchronous, exactly what FheTaskGpu::run returns, cxx_fhe_task_gpu.cpp:110-116)
    api = None
    if not args.no_api:
        rlk_host = rlk_dev.cpu()
        out_host.fill_(-1)  # sentinel: the API arm must overwrite every output limb itself
        arm = RunnerArm(P, B, a_host, b_host, rlk_host, out_host, local, "r%d" % rank)
        try:
            arm.step()  # warm-up: context, tables, pools, key upload
            arm.step()
            api_steps = max(1, min(args.steps, 5))
            barrier()
            t0 = time.perf_counter()
            for _ in range(api_steps):
                arm.step()
            torch.cuda.synchronize()
            dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            dt = float(dt.item())
            st = arm.stats()
            api = {"value": world * B * api_steps / dt, "unit": UNIT, "ms_per_step": 1e3 * dt / api_steps,
                   "h2d_bytes_per_step": (a_host.numel() + b_host.numel()) * 8, "d2h_bytes_per_step": out_host.numel() * 8,
                   "path": "create_/bind_/run_fhe_gpu_task (include/lattisense_gpu_runner.h), %d-node mega_ag graph, "
                           "test plug-in EXPORT/IMPORT over host limb buffers" % (3 * B),
                   "runner_stats": {"kernel_launches": st[0], "batched_groups": st[1], "fused_nodes": st[2]},
                   "host_profile_last_step": arm.host_profile(),
                   "matches_device_path": bool(torch.equal(out_host, out_dev.cpu()))}
        finally:
            arm.close()

    # roofline pass: per-kernel-class CUDA events on the launching stream (one extra step, same inputs)
    ctx.profile(True)
    step_dev()
    torch.cuda.synchronize()
    kinds = ["ntt_fwd", "ntt_inv", "basis_ext", "ks_inner", "elementwise"]
    prof = {k: ctx.profile_read(i) for i, k in enumerate(kinds)}
    ctx.profile(False)
    peak, peak_src = peaks()
    try:
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
            traffic = json.load(f)
    except Exception:
        traffic = {}
    tot_ms = sum(v[0] for v in prof.values())
    ntt_ms = prof["ntt_fwd"][0] + prof["ntt_inv"][0]
    ntt_bytes = prof["ntt_fwd"][2] + prof["ntt_inv"][2]
    ntt_k = prof["ntt_fwd"][1] + prof["ntt_inv"][1]
    achieved = ntt_bytes / (ntt_ms / 1e3) / 1e9 if ntt_ms > 0 else 0.0

    # stand-alone NTT throughput: 8*B limb transforms in one launch, in place
    mi = (C.c_int * l)(*range(l))
    for _ in range(3):
        L._check(lib.lsb_ntt(ctx.h, a_dev.data_ptr(), a_dev.data_ptr(), 2 * l, mi, l, 0, B, stream))
    reps = 10
    ms_ntt = timed(lambda: L._check(lib.lsb_ntt(ctx.h, a_dev.data_ptr(), a_dev.data_ptr(), 2 * l, mi, l, 0, B,
                                                stream)), reps)
    ntt_alone = 16.0 * N * 2 * l * B * reps / (ms_ntt / 1e3) / 1e9

    # the rest of BASELINE.json's metric ("mult_relin + rotate ops/s, N=2^14..2^16"): device-resident side numbers,
    # same timing rules, single GPU only (parity of these paths: tests/test_gpu_parity.py, tests/test_runner_gpu.py)
    secondary = None
    if world == 1 and not args.no_secondary:
        secondary = {}

        def rate(fn, units, steps=5):
            for _ in range(3):
                fn()
            return units * steps / (timed(fn, steps) / 1e3)

        # CKKS rotate (automorphism + key-switch), same ring / level / batch as the headline
        gal = R.galois_for_col_step(N, 1)
        glk_dev = rand_limbs((P.beta(LEVEL), 2), P.q[:l] + P.p)
        rot_out = torch.empty((B, 2, l, N), dtype=torch.int64, device=dev)
        secondary["ckks_rotate_n16384_level3_ops_per_s"] = rate(lambda: L._check(lib.lsb_rotate(
            ctx.h, LEVEL, a_dev.data_ptr(), glk_dev.data_ptr(), LEVEL, gal, rot_out.data_ptr(), B, stream)), B)
        del glk_dev, rot_out
        # BFV mult_relin (configs[0]: n=16384, t=65537, level 2): 54-57-bit primes -> integer NTT policy
        PB, lb, Bb = R.BFV_N14, 3, 256
        cb = L.Context(PB.scheme, PB.logN, PB.q, PB.p, PB.t, device=local)
        cb.stream = stream

        def limbs(lead, mods, n):
            t = torch.empty(lead + (len(mods), n), dtype=torch.int64, device=dev)
            for i, m in enumerate(mods):
                t[..., i, :] = torch.randint(0, m, lead + (n,), dtype=torch.int64, device=dev)
            return t

        xa, xb = limbs((Bb, 2), PB.q[:lb], PB.N), limbs((Bb, 2), PB.q[:lb], PB.N)
        rk = limbs((PB.beta(lb - 1), 2), PB.q[:lb] + PB.p, PB.N)
        xo = torch.empty((Bb, 2, lb, PB.N), dtype=torch.int64, device=dev)
        secondary["bfv_mult_relin_n16384_level2_ops_per_s"] = rate(lambda: L._check(lib.lsb_mult_relin(
            cb.h, lb - 1, xa.data_ptr(), xb.data_ptr(), rk.data_ptr(), lb - 1, xo.data_ptr(), Bb, stream)), Bb)
        del xa, xb, rk, xo
        cb.close()
        # CKKS N=2^16 (bootstrap chain: 61-bit and 40-bit primes, k=5): rotations at level 2 (configs[3]) and one
        # full-depth relinearisation key-switch (level 24, beta=5: the unit configs[4] is made of)
        P16 = R.CKKS_BTP_N16
        c16 = L.Context(P16.scheme, P16.logN, P16.q, P16.p, 0, device=local)
        c16.stream = stream
        l16, B16 = 3, 64
        ct = limbs((B16, 2), P16.q[:l16], P16.N)
        gk = limbs((P16.beta(l16 - 1), 2), P16.q[:l16] + P16.p, P16.N)
        ro = torch.empty_like(ct)
        g16 = R.galois_for_col_step(P16.N, 1)
        secondary["ckks_rotate_n65536_level2_ops_per_s"] = rate(lambda: L._check(lib.lsb_rotate(
            c16.h, l16 - 1, ct.data_ptr(), gk.data_ptr(), l16 - 1, g16, ro.data_ptr(), B16, stream)), B16)
        del ct, gk, ro
        l25, B25 = 25, 8
        c3 = limbs((B25, 3), P16.q[:l25], P16.N)
        rk = limbs((P16.beta(l25 - 1), 2), P16.q[:l25] + P16.p, P16.N)
        o2 = torch.empty((B25, 2, l25, P16.N), dtype=torch.int64, device=dev)
        secondary["ckks_keyswitch_n65536_level24_ops_per_s"] = rate(lambda: L._check(lib.lsb_relin(
            c16.h, l25 - 1, c3.data_ptr(), rk.data_ptr(), l25 - 1, o2.data_ptr(), B25, stream)), B25)
        del c3, rk, o2
        c16.close()
        # the reference frontend's own default CKKS sets for n = 32768 and n = 65536 (frontend/parameter.json) at their top
        # level: mult_relin_rescale and rotate ("N = 2^14..2^16" of BASELINE.json's metric; 41/46-bit Q primes on the FP64
        # policy, 51/56-bit q0 and P primes on the integer policy; parity: tests/test_gpu_parity.py
        # test_keyswitch_with_more_digits_than_one_launch_group)
        for Pd, Bd in ((R.CKKS_N15, 64), (R.CKKS_N16, 16)):
            try:
                cd = L.Context(Pd.scheme, Pd.logN, Pd.q, Pd.p, 0, device=local)
                cd.stream = stream
                ld = len(Pd.q)
                xa, xb = limbs((Bd, 2), Pd.q, Pd.N), limbs((Bd, 2), Pd.q, Pd.N)
                kd = limbs((Pd.beta(ld - 1), 2), Pd.q + Pd.p, Pd.N)
                xo = torch.empty((Bd, 2, ld - 1, Pd.N), dtype=torch.int64, device=dev)
                secondary["ckks_mult_relin_rescale_n%d_level%d_ops_per_s" % (Pd.N, ld - 1)] = rate(lambda: L._check(
                    lib.lsb_mult_relin_rescale(cd.h, ld - 1, xa.data_ptr(), xb.data_ptr(), kd.data_ptr(), ld - 1, xo.data_ptr(), Bd,
                                               stream)), Bd, steps=3)
                gd = R.galois_for_col_step(Pd.N, 1)
                secondary["ckks_rotate_n%d_level%d_ops_per_s" % (Pd.N, ld - 1)] = rate(lambda: L._check(lib.lsb_rotate(
                    cd.h, ld - 1, xa.data_ptr(), kd.data_ptr(), ld - 1, gd, xb.data_ptr(), Bd, stream)), Bd, steps=3)
                del xa, xb, kd, xo
                cd.close()
            except Exception as ex:  # a side measurement never costs the headline line
                secondary["ckks_n%d_error" % Pd.N] = str(ex)[:200]
        # CKKS bootstrapping N=2^16 (configs[4]): one `bootstrap` node = level 0 -> level 9 on the 25 Q + 5 P chain; random
        # residues for ciphertext and keys (the op sequence is data-independent; parity / precision: tests/test_bootstrap_gpu.py,
        # tests/test_runner_gpu.py[ckks_btp64k_bootstrap])
        try:
            import btp_ref as BT
            spec = BT.BtpSpec(P16)
            c16 = L.Context(P16.scheme, P16.logN, P16.q, P16.p, 0, device=local)
            c16.stream = stream
            bt = L.Bootstrapper(c16, list(spec.ints()), list(spec.doubles()))
            rots = bt.rotations()
            top = len(P16.q) - 1
            # 4 x 150 MiB stand in for the 51 keys (same traffic per use)
            key_pool = [limbs((P16.beta(top), 2), P16.q + P16.p, P16.N) for _ in range(4)]
            gals = [R.galois_for_col_step(P16.N, r) for r in rots] + [R.galois_for_row(P16.N)]
            garr = np.array(gals, dtype=np.uint64)
            gptr = (C.c_void_p * len(gals))(*[key_pool[i % 4].data_ptr() for i in range(len(gals))])
            dts = limbs((1, 2), [P16.q[0]] + P16.p, P16.N)
            for Bb_ in (1, 4):
                ct0 = limbs((Bb_, 2), P16.q[:1], P16.N)
                outb = torch.empty((Bb_, 2, spec.out_level + 1, P16.N), dtype=torch.int64, device=dev)
                stats = (C.c_longlong * 5)()
                run = lambda: L._check(lib.lsb_bootstrap_run(bt.h, ct0.data_ptr(), key_pool[0].data_ptr(), top, garr.ctypes.data_as(L.u64p),
                                                             gptr, len(gals), dts.data_ptr(), key_pool[1].data_ptr(), outb.data_ptr(), Bb_,
                                                             stream, stats))
                run()  # encodes the CoeffsToSlots / SlotsToCoeffs plaintexts once
                torch.cuda.synchronize()
                ms_b = timed(run, 2) / 2
                secondary["ckks_bootstrap_n65536_per_s_batch%d" % Bb_] = Bb_ / (ms_b / 1e3)
                secondary["ckks_bootstrap_n65536_ms_batch%d" % Bb_] = ms_b
            secondary["ckks_bootstrap_key_switches"] = int(stats[1])
            secondary["ckks_bootstrap_ops"] = {"rotations": int(stats[0]), "ct_x_ct": int(stats[2]), "rescales": int(stats[3]), "ct_x_pt": int(stats[4])}
            bt.close()
            c16.close()
        except Exception as ex:  # never lose the headline line to a side measurement
            secondary["ckks_bootstrap_error"] = str(ex)[:200]
        if not args.no_api:
            secondary["reference_benchmark_gpu"] = reference_benchmark_gpu(local)
        secondary["note"] = ("device-resident, CUDA events, batches 1024 / 256 / 64 / 8 (default n=32768 / 65536 sets: 64 / 16); random residues (throughput does "
                             "not depend on values)")

    # multi-GPU only: the path's one exchange step -- the limb-sharded key-switch for N=2^16 through
    # the library's C++ NCCL driver (lsb_ks_shard_*), batch 16, every rank's limbs checked against the oracle
    limb = None
    if world > 1 and not args.no_secondary:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import ks_shard_nccl_run as KS
            from lattisense_b200.dist import Ranks
            limb = KS.measure(16, 24, 16, 10, check=True, rk=Ranks())
        except Exception as ex:
            limb = {"error": str(ex)[:300]}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    cpu = None
    if world == 1 and not args.no_cpu:
        per_core = 800  # about 10 s of CPU work on every core
        v, dt = cpu_ops_per_s(cores, per_core)
        v_simple, _ = cpu_ops_per_s(cores, 40, "oracle")
        cpu = {"value": v, "unit": UNIT, "cores": min(cores, 32), "kind": "port", "variant": "optimised",
               "oracle_variant_value": v_simple,
               "sample": "%d mult_relin_rescale ops (%d per core, %.1f s) of the same workload; optimised CPU port "
                         "(oracle/orc_fast.hpp), simple oracle beside it" % (per_core * min(cores, 32), per_core, dt)}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "CKKS mult_relin+rescale n=16384 level 3->2 (BASELINE configs[1])",
                   "batch_per_gpu": B, "l2": "inputs (2 GiB/step) larger than L2, no flush needed",
                   "parallelism": "batch-sharded x%d, no collective" % world},
        "e2e": api if api is not None else None,
        "e2e_kernel_abi": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": (a_host.numel() + b_host.numel()) * 8,
                           "d2h_bytes_per_step": out_host.numel() * 8, "ms_per_step": ms_e2e / e2e_steps,
                           "path": "C ABI lsb_upload + lsb_mult_relin_rescale + lsb_download (include/lsb200.h), pinned host buffers"},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "k_ntt_cluster<14,*,ArithF64> (forward incl. its fused finish epilogue + inverse, in-step average)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     # DRAM bytes per NTT launch and the step-level traffic ratio come from an ncu capture of one step of this
                     # build (tools/step_traffic.py -> profiles/r02_traffic.json), NOT from the timed run: labelled as such
                     "traffic": traffic.get("ntt_dram_bytes_per_launch"), "traffic_source": traffic.get("source"),
                     "step_traffic_ratio": traffic.get("step_traffic_ratio"), "step_dram_bytes": traffic.get("step_dram_bytes"),
                     "step_fp64_pipe_busy_ms": traffic.get("fp64_pipe_busy_ms"),
                     "algorithmic_bytes_per_launch": ntt_bytes / max(ntt_k, 1),
                     # second roof: the FP64 pipe, 1.015 M DFMA-class ops per limb transform at the measured
                     # 57.8 lanes/clk/SM = 4.36 TB/s of algorithmic NTT traffic if the pipe never idled
                     "fp64_pipe_roof_gbs": 4360.0, "frac_of_fp64_roof": achieved / 4360.0,
                     "peak_source": peak_src, "launches": ntt_k, "ms_per_launch": ntt_ms / max(ntt_k, 1),
                     "share_of_step": ntt_ms / tot_ms if tot_ms else None,
                     "standalone_ntt_gbs": ntt_alone, "standalone_frac": ntt_alone / peak},
        "kernel_ms_per_step": {k: v[0] for k, v in prof.items()},
        "secondary": secondary,
        "limb_sharded_keyswitch": limb,
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-api", action="store_true", help="skip the run_fhe_gpu_task end-to-end leg")
    ap.add_argument("--no-secondary", action="store_true", help="skip the rotate / BFV / N=2^16 side measurements")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
