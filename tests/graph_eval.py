"""Drives a mega_ag task directory through the C-ABI runner (create_/bind_/run_/release_fhe_gpu_task via
the test plug-in tests/harness/harness.cpp) and evaluates the same graph with the CPU oracle.

TEST INFRASTRUCTURE: the oracle evaluation here is the checker; the product path is the runner library.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

import fhe_ref as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GRAPHS = os.path.join(ROOT, "tests", "golden", "graphs")
LIB_DIR = os.path.join(ROOT, "lattisense_b200", "_lib")
HARNESS_SO = os.path.join(ROOT, "tests", "_build", "libharness.so")

TYPE_PLAINTEXT, TYPE_CIPHERTEXT, TYPE_RELIN_KEY, TYPE_GALOIS_KEY, TYPE_SWITCH_KEY, TYPE_CUSTOM = range(6)


class THandle(C.Structure):
    pass


THandle._fields_ = [("kind", C.c_int), ("level", C.c_int), ("degree", C.c_int), ("n", C.c_int), ("k", C.c_int),
                    ("data", C.POINTER(C.c_uint64)), ("n_keys", C.c_int), ("galois_elements", C.POINTER(C.c_uint64)),
                    ("keys", C.POINTER(C.POINTER(THandle)))]

_h = None


def harness():
    """Builds (if stale) and loads the test plug-in; it links against the product library."""
    global _h
    if _h is not None:
        return _h
    src = os.path.join(ROOT, "tests", "harness", "harness.cpp")
    lib = os.path.join(LIB_DIR, "libgpu_mega_ag_runner.so")
    if not os.path.exists(lib):
        raise RuntimeError("libgpu_mega_ag_runner.so is missing: run __graft_entry__.build() (no CPU fallback)")
    deps = [src, lib] + [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
    if not os.path.exists(HARNESS_SO) or os.path.getmtime(HARNESS_SO) < max(os.path.getmtime(p) for p in deps):
        os.makedirs(os.path.dirname(HARNESS_SO), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "include"),
                               "-o", HARNESS_SO, src, "-L" + LIB_DIR, "-lgpu_mega_ag_runner",
                               "-Wl,-rpath," + LIB_DIR])
    h = C.CDLL(HARNESS_SO)
    h.th_open.restype = C.c_void_p
    h.th_open.argtypes = [C.c_char_p, C.c_int]
    h.th_close.argtypes = [C.c_void_p]
    h.th_last_error.restype = C.c_char_p
    ip = C.POINTER(C.c_int)
    hp = C.POINTER(C.POINTER(THandle))
    h.th_run.argtypes = [C.c_void_p, C.c_int, ip, ip, ip, hp, C.c_int, ip, ip, ip, hp, C.c_int, ip]
    _h = h
    return h


def runner_lib():
    L = C.CDLL(os.path.join(LIB_DIR, "libgpu_mega_ag_runner.so"))
    L.lsb_gpu_task_describe.argtypes = [C.c_void_p, C.POINTER(C.c_longlong)]
    L.lsb_gpu_task_shard_preview.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.c_int]
    L.lsb_gpu_task_set_devices.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.c_int]
    L.lsb_gpu_task_last_stats.argtypes = [C.c_void_p] + [C.POINTER(C.c_longlong)] * 3
    L.lsb_gpu_task_transfer_stats.argtypes = [C.c_void_p] + [C.POINTER(C.c_longlong)] * 4
    L.lsb_gpu_task_set_key_cache.argtypes = [C.c_void_p, C.c_int]
    L.lsb_gpu_task_invalidate_keys.argtypes = [C.c_void_p]
    return L


def params_of_graph(g, name):
    p = g["parameter"]
    scheme = R.SCHEME_BFV if g["algorithm"] == "BFV" else R.SCHEME_CKKS
    logn = p["n"].bit_length() - 1
    nq = min(len(p["q"]), p.get("max_level", len(p["q"]) - 1) + 1)
    return R.Params(name, scheme, logn, [int(x) for x in p["q"][:nq]], [int(x) for x in p["p"]],
                    int(p.get("t", 0)) if scheme == R.SCHEME_BFV else 0, float(p.get("scale", 0.0)))


def _mulmod_rows(x, y, q):
    return ((x.astype(object) * y.astype(object)) % q).astype(np.uint64)


class GraphCase:
    """One task directory: random inputs (seeded), oracle evaluation, run through the runner."""

    def __init__(self, path, seed=7, pt_pool=0):
        """pt_pool > 0: plaintext inputs of one level share pt_pool distinct random arrays (large graphs)."""
        self.path = path
        self.pt_pool = pt_pool
        with open(os.path.join(path, "mega_ag.json")) as f:
            self.g = json.load(f)
        with open(os.path.join(path, "task_signature.json")) as f:
            self.sig = json.load(f)
        self.params = params_of_graph(self.g, os.path.basename(path))
        self.o = R.Oracle(self.params)
        self.N = self.params.N
        self.rng = np.random.default_rng(seed)
        self.data = {int(k): v for k, v in self.g["data"].items()}
        self.compute = {int(k): v for k, v in self.g["compute"].items()}
        self.values = {}
        self.secret = self.o.gen_secret(seed)
        self.btp = None
        if any(c["type"] == "bootstrap" for c in self.compute.values()):
            # bootstrap graphs need real key material and real encryptions: dense secret of weight 192, ephemeral
            # sparse secret of weight 32 behind swk_dts / swk_std (what CkksBtpContext generates on the reference side)
            import btp_ref as B
            p = self.g["parameter"]
            self.btp = B.BtpSpec(self.params, p["btp_cts_start_level"], p["btp_cts_depth"], p["btp_eval_mod_start_level"],
                                 p["btp_eval_mod_k"], p["btp_eval_mod_sine_deg"], p["btp_eval_mod_double_angle"],
                                 p["btp_stc_start_level"], p["btp_stc_depth"], p["btp_output_level"],
                                 p["btp_eval_mod_message_ratio"], p["btp_cts_bsgs_ratio"])
            self.secret = B.fixed_weight_secret(self.N, 192, seed)
            self.secret_sparse = B.fixed_weight_secret(self.N, 32, seed + 1)
            self.messages = {}
        self._make_inputs(seed)

    # ---- inputs ----
    def _rand_limbs(self, lead, mods):
        out = np.zeros(tuple(lead) + (len(mods), self.N), dtype=np.uint64)
        for i, m in enumerate(mods):
            out[..., i, :] = self.rng.integers(0, m, size=tuple(lead) + (self.N,), dtype=np.uint64)
        return out

    def _make_inputs(self, seed):
        q, p = self.params.q, self.params.p
        self.glk = {}
        self._pool = {}
        for idx in self.g["inputs"]:
            d = self.data[idx]
            if d.get("is_custom"):
                # custom payload consumed by an `encode_ringt` / `encode` custom node: generated in the shape of the
                # plaintext that node outputs (the test plug-in's encoders are identity, tests/harness/harness.cpp)
                cons = [c for c in self.compute.values() if idx in c["inputs"]][0]
                od = self.data[cons["outputs"][0]]
                if od["type"] == "pt_ringt":
                    tt = self.params.t if self.params.scheme == R.SCHEME_BFV else 65537
                    self.values[idx] = self.rng.integers(0, tt, size=(1, self.N), dtype=np.uint64)
                else:
                    self.values[idx] = self._rand_limbs((), q[:od["level"] + 1])
                continue
            t, lv = d["type"], d["level"]
            if t in ("ct", "ct3") and self.btp is not None:
                z = self.rng.uniform(-1, 1, self.N // 2) + 1j * self.rng.uniform(-1, 1, self.N // 2)
                self.messages[idx] = z
                self.values[idx] = self.o.encrypt(self.secret, seed + 100 + idx, lv, R.ckks_encode(self.o, z, lv, self.params.scale))
            elif t in ("ct", "ct3"):
                self.values[idx] = self._rand_limbs((d["degree"] + 1,), q[:lv + 1])
            elif t in ("pt", "pt_mul"):
                if self.pt_pool:
                    key = (lv, idx % self.pt_pool)
                    if key not in self._pool:
                        self._pool[key] = self._rand_limbs((), q[:lv + 1])
                    self.values[idx] = self._pool[key]
                else:
                    self.values[idx] = self._rand_limbs((), q[:lv + 1])
            elif t == "pt_ringt":
                tt = self.params.t if self.params.scheme == R.SCHEME_BFV else 65537
                self.values[idx] = self.rng.integers(0, tt, size=(1, self.N), dtype=np.uint64)
            elif t == "swk" and self.btp is not None:
                import btp_ref as B
                evk = np.zeros(self.o.evk_shape(lv), dtype=np.uint64)
                s_in, s_out = (self.secret, self.secret_sparse) if d["id"] == "swk_dts" else (self.secret_sparse, self.secret)
                B.lib().orc_gen_swk(self.o.hp, s_in.ctypes.data_as(B.i8p), s_out.ctypes.data_as(B.i8p), C.c_uint64(seed + 4 + lv), lv, R.P(evk))
                self.values[idx] = evk
            elif t in ("rlk", "swk"):
                # a real key for the task's secret at the node's level (uniform limbs would do for parity, but
                # real keys let the decode-level checks run on the same data)
                self.values[idx] = self.o.gen_rlk(self.secret, seed + 1, lv)
            elif t == "glk":
                gal = int(d["galois_element"])
                self.values[idx] = self.o.gen_glk(self.secret, seed + 2 + gal % 1000, lv, gal)
                self.glk[gal] = idx
            else:
                raise ValueError("unsupported input type " + t)

    # ---- oracle evaluation of the graph ----
    def _lift_ringt(self, m, level):
        l = level + 1
        return self.o.ntt_limbs(np.repeat(m, l, axis=0))

    def _mul_plain(self, ct, pt_ntt):
        o, bfv = self.o, self.params.scheme == R.SCHEME_BFV
        x = o.ntt_limbs(ct) if bfv else ct
        out = np.zeros_like(x)
        for i in range(x.shape[1]):
            for pp in range(x.shape[0]):
                out[pp, i] = _mulmod_rows(x[pp, i], pt_ntt[i], self.params.q[i])
        return out  # NTT domain

    def _to_native(self, x_ntt):
        return self.o.ntt_limbs(x_ntt, inverse=True) if self.params.scheme == R.SCHEME_BFV else x_ntt

    def _eval_node(self, c):
        o, D, V = self.o, self.data, self.values
        ins = c["inputs"]
        t = c["type"]
        bfv = self.params.scheme == R.SCHEME_BFV
        kind = lambda i: D[i]["type"]
        if t in ("add", "sub"):
            a, b = ins if len(ins) == 2 else (ins[0], ins[0])
            fn = o.add if t == "add" else o.sub
            if kind(a) in ("ct", "ct3") and kind(b) in ("ct", "ct3"):
                return fn(V[a], V[b])
            ci, pi = (a, b) if kind(a) in ("ct", "ct3") else (b, a)
            ct = V[ci]
            lv = D[ci]["level"]
            if kind(pi) == "pt_ringt":
                pt = o.bfv_scale_up(lv, V[pi][0]) if bfv else self._lift_ringt(V[pi], lv)
            else:
                pt = V[pi]
            out = ct.copy()
            if ci == a:
                out[0] = fn(ct[:1], pt[None])[0]
            else:  # pt (op) ct
                out = ct.copy() if t == "add" else o.neg(ct)
                out[0] = fn(pt[None], ct[:1])[0]
            return out
        if t == "neg":
            return o.neg(V[ins[0]])
        if t == "mult":
            a, b = ins if len(ins) == 2 else (ins[0], ins[0])
            if kind(a) in ("ct", "ct3") and kind(b) in ("ct", "ct3"):
                return o.mult(V[a], V[b])
            ci, pi = (a, b) if kind(a) in ("ct", "ct3") else (b, a)
            lv = D[ci]["level"]
            pt = self._lift_ringt(V[pi], lv) if kind(pi) == "pt_ringt" else V[pi]
            return self._to_native(self._mul_plain(V[ci], pt))
        if t == "relin":
            return o.relin(V[ins[0]], V[ins[1]])
        if t == "rescale":
            return o.rescale(V[ins[0]])
        if t == "drop_level":
            return o.drop_level(V[ins[0]])
        if t in ("rotate_col", "rotate_row"):
            gal = int(D[ins[1]]["galois_element"])
            return o.rotate(V[ins[0]], V[ins[1]], gal)
        if t in ("cmp_sum", "cmpac_sum"):
            n = c["sum_cnt"]
            acc_in = V[ins[n]] if t == "cmpac_sum" else None
            pts = ins[n + (1 if acc_in is not None else 0):]
            lv = D[ins[0]]["level"]
            acc = None
            for i in range(n):
                pt = self._lift_ringt(V[pts[i]], lv) if kind(pts[i]) == "pt_ringt" else V[pts[i]]
                term = self._mul_plain(V[ins[i]], pt)
                acc = term if acc is None else o.add(acc, term)
            out = self._to_native(acc)
            return o.add(out, acc_in) if acc_in is not None else out
        if t == "bootstrap":
            import btp_ref as B

            class K:
                pass
            k = K()
            k.rlk = V[ins[1]]
            k.top = D[ins[1]]["level"]
            glk_ids = [i for i in ins[2:] if kind(i) == "glk"]
            k.gals = [int(D[i]["galois_element"]) for i in glk_ids]
            k.glk = [V[i] for i in glk_ids]
            swk = [i for i in ins if kind(i) == "swk"]
            k.swk_dts, k.swk_std = V[swk[0]], V[swk[1]]
            out, self.btp_stats = B.oracle_bootstrap(o, self.btp, k, V[ins[0]][None])
            return out[0]
        if c.get("is_custom"):
            if t in ("encode_ringt", "encode"):
                return V[ins[0]]
            if t == "passthrough":
                return V[ins[0]]
            raise ValueError("unknown custom node " + t)
        raise ValueError("unsupported node type " + t)

    def expected(self, free=False):
        """free=True drops an intermediate once its last consumer has run (large graphs)."""
        pending = dict(self.compute)
        uses = {}
        for c in self.compute.values():
            for i in c["inputs"]:
                uses[i] = uses.get(i, 0) + 1
        keep = set(self.g["inputs"]) | set(self.g["outputs"])
        while pending:
            progressed = False
            for k in sorted(pending):
                c = pending[k]
                if all(i in self.values for i in c["inputs"]):
                    self.values[c["outputs"][0]] = self._eval_node(c)
                    del pending[k]
                    progressed = True
                    if free:
                        for i in c["inputs"]:
                            uses[i] -= 1
                            if uses[i] == 0 and i not in keep:
                                del self.values[i]
            if not progressed:
                raise RuntimeError("graph evaluation stuck")
        return [self.values[i] for i in self.g["outputs"]]

    # ---- handles / arguments ----
    def _handle(self, idx, arr, keep):
        d = self.data[idx]
        t = d["type"]
        h = THandle()
        arr = np.ascontiguousarray(arr, dtype=np.uint64)
        keep.append(arr)
        h.n, h.k = self.N, len(self.params.p)
        h.data = arr.ctypes.data_as(C.POINTER(C.c_uint64))
        if d.get("is_custom"):
            h.level, h.degree, h.kind = arr.shape[0] - 1, 0, 5
            return h
        h.level, h.degree = d["level"], d.get("degree", 0)
        h.kind = {"ct": 0, "ct3": 0, "pt": 1, "pt_mul": 1, "pt_ringt": 2, "rlk": 3, "swk": 3, "glk": 3}[t]
        return h

    def build_args(self):
        """CArgument lists in the order FheTaskGpu builds them: data arguments from task_signature.json's `online`
        list (each `size` handles), then rlk, then ONE glk handle (cxx_sdk_v2/cxx_argument.h:178-260)."""
        keep = []
        in_ids = list(self.g["inputs"])
        data_ids = [i for i in in_ids if self.data[i].get("is_custom") or self.data[i]["type"] not in ("rlk", "glk", "swk")]
        args_in, handles_in = [], []
        pos = 0
        for a in self.sig["online"]:
            if a["phase"] != "in":
                continue
            cnt = int(np.prod(a["size"]))
            ids = data_ids[pos:pos + cnt]
            pos += cnt
            is_custom = any(self.data[i].get("is_custom") for i in ids)
            ty = TYPE_CUSTOM if is_custom else TYPE_CIPHERTEXT if a["type"] in ("ct", "ct3") else TYPE_PLAINTEXT
            args_in.append((ty, a.get("level", 0), cnt))
            for i in ids:
                handles_in.append(self._handle(i, self.values[i], keep))
        assert pos == len(data_ids), "task_signature does not cover the data inputs"
        for i in in_ids:
            if not self.data[i].get("is_custom") and self.data[i]["type"] == "rlk":
                args_in.append((TYPE_RELIN_KEY, self.data[i]["level"], 1))
                handles_in.append(self._handle(i, self.values[i], keep))
        if self.glk:
            gals = sorted(self.glk)
            ga = np.array(gals, dtype=np.uint64)
            keep.append(ga)
            subs = [self._handle(self.glk[g_], self.values[self.glk[g_]], keep) for g_ in gals]
            for s in subs:
                s.kind = 3
            keep.append(subs)
            parr = (C.POINTER(THandle) * len(subs))(*[C.pointer(s) for s in subs])
            keep.append(parr)
            hs = THandle()
            hs.kind, hs.n, hs.k, hs.n_keys = 4, self.N, len(self.params.p), len(subs)
            hs.galois_elements = ga.ctypes.data_as(C.POINTER(C.c_uint64))
            hs.keys = parr
            args_in.append((TYPE_GALOIS_KEY, -1, 1))
            handles_in.append(hs)
        for i in in_ids:
            if not self.data[i].get("is_custom") and self.data[i]["type"] == "swk":
                args_in.append((TYPE_SWITCH_KEY, self.data[i]["level"], 1))
                handles_in.append(self._handle(i, self.values[i], keep))
        args_out, handles_out, outs = [], [], []
        for i in self.g["outputs"]:
            d = self.data[i]
            arr = np.zeros((d["degree"] + 1, d["level"] + 1, self.N), dtype=np.uint64)
            outs.append(arr)
            args_out.append((TYPE_CIPHERTEXT, d["level"], 1))
            handles_out.append(self._handle(i, arr, keep))
        return args_in, handles_in, args_out, handles_out, outs, keep

    def run(self, device=0, task=None, devices=None, reuse_args=False):
        """reuse_args: keep the argument handles of the previous run (stable key handles: what a caller that holds its keys
        across runs passes, and what the runner's key cache is keyed by)"""
        h = harness()
        own = task is None
        if own:
            task = h.th_open(self.path.encode(), 1)
            if not task:
                raise RuntimeError("th_open: " + h.th_last_error().decode())
        try:
            if devices:
                runner_lib().lsb_gpu_task_set_devices(task, (C.c_int * len(devices))(*devices), len(devices))
            if not (reuse_args and getattr(self, "_args", None)):
                self._args = self.build_args()
            args_in, handles_in, args_out, handles_out, outs, keep = self._args
            for o in outs:
                o[...] = 0

            def pack(args, handles):
                ty = (C.c_int * len(args))(*[a[0] for a in args])
                lv = (C.c_int * len(args))(*[a[1] for a in args])
                sz = (C.c_int * len(args))(*[a[2] for a in args])
                hp = (C.POINTER(THandle) * len(handles))(*[C.pointer(x) for x in handles])
                return ty, lv, sz, hp

            ity, ilv, isz, ihp = pack(args_in, handles_in)
            oty, olv, osz, ohp = pack(args_out, handles_out)
            prog = (C.c_int * 3)()
            rc = h.th_run(task, len(args_in), ity, ilv, isz, ihp, len(args_out), oty, olv, osz, ohp, device, prog)
            if rc != 0:
                raise RuntimeError("run_fhe_gpu_task failed: " + h.th_last_error().decode())
            self.progress = tuple(prog)
            st = [C.c_longlong() for _ in range(3)]
            runner_lib().lsb_gpu_task_last_stats(task, *[C.byref(x) for x in st])
            self.stats = tuple(x.value for x in st)
            tr = [C.c_longlong() for _ in range(4)]
            runner_lib().lsb_gpu_task_transfer_stats(task, *[C.byref(x) for x in tr])
            self.transfer = dict(zip(["h2d", "d2h", "zero_copy", "key_cached"], [x.value for x in tr]))
            return outs
        finally:
            if own:
                h.th_close(task)


def all_cases():
    return sorted(d for d in os.listdir(GRAPHS) if os.path.exists(os.path.join(GRAPHS, d, "mega_ag.json")))


def unpack_case(case, dest):
    """Large fixtures are committed gzipped (tests/golden/graphs/<case>/*.json.gz); the runner wants plain files."""
    import gzip
    import shutil
    os.makedirs(dest, exist_ok=True)
    for f in ("mega_ag.json", "task_signature.json"):
        with gzip.open(os.path.join(GRAPHS, case, f + ".gz"), "rb") as src, open(os.path.join(dest, f), "wb") as dst:
            shutil.copyfileobj(src, dst)
    return dest


def digest(arrays):
    import hashlib
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a, dtype=np.uint64).tobytes())
    return h.hexdigest()
