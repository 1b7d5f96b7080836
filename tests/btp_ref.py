"""Test support for the bootstrap node: key material, the ORACLE run of the product's pipeline
(tests/btp/btp_oracle.cpp = lattisense_b200/csrc/lsb_bootstrap.h instantiated with the CPU oracle's ciphertext operations)
and the reference's acceptance metric (decrypt + decode, mean real precision, unittests/test_gpu_ckks.cpp:37-43, 743-808).
TEST INFRASTRUCTURE: never imported by the product package."""
import ctypes as C
import os
import subprocess

import numpy as np

import fhe_ref as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "tests", "_build", "libbtp_orc.so")
SRC = os.path.join(ROOT, "tests", "btp", "btp_oracle.cpp")
u64p = C.POINTER(C.c_uint64)
i8p = C.POINTER(C.c_int8)

_lib = None


def lib():
    global _lib
    if _lib is None:
        deps = [SRC, os.path.join(ROOT, "lattisense_b200", "csrc", "lsb_bootstrap.h")] + \
               [os.path.join(ROOT, "oracle", f) for f in ("orc_core.hpp", "orc_keys.hpp", "orc_math.hpp")]
        if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(p) for p in deps):
            os.makedirs(os.path.dirname(SO), exist_ok=True)
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-o", SO, SRC])
        _lib = C.CDLL(SO)
    return _lib


class BtpSpec:
    """bootstrapping parameters as mega_ag.json carries them (frontend/custom_task.py:424-463 defaults)"""

    def __init__(self, params, cts_start=24, cts_depth=4, evalmod_start=20, K=16, sine_deg=30, double_angle=3, stc_start=12,
                 stc_depth=3, out_level=9, message_ratio=256.0, ratio=2.0):
        self.params = params
        self.log_slots = params.logN - 1
        self.cts_start, self.cts_depth, self.evalmod_start = cts_start, cts_depth, evalmod_start
        self.K, self.sine_deg, self.double_angle = K, sine_deg, double_angle
        self.stc_start, self.stc_depth, self.out_level = stc_start, stc_depth, out_level
        self.message_ratio, self.ratio = message_ratio, ratio

    def ints(self):
        return (C.c_int * 11)(self.params.logN, self.log_slots, self.cts_start, self.cts_depth, self.evalmod_start, self.K,
                              self.sine_deg, self.double_angle, self.stc_start, self.stc_depth, self.out_level)

    def doubles(self):
        return (C.c_double * 4)(self.params.scale, self.ratio, self.message_ratio, self.ratio)

    def json_fields(self):
        p = self.params
        return {"n": p.N, "max_level": len(p.q) - 1, "q": p.q, "p": p.p, "slots": p.N // 2, "scale": p.scale,
                "btp_cts_start_level": self.cts_start, "btp_cts_depth": self.cts_depth, "btp_cts_bsgs_ratio": self.ratio,
                "btp_eval_mod_q": p.q[0], "btp_eval_mod_start_level": self.evalmod_start,
                "btp_eval_mod_scaling_factor": 1 << 60, "btp_eval_mod_sine_type": "Cos1",
                "btp_eval_mod_message_ratio": self.message_ratio, "btp_eval_mod_k": self.K,
                "btp_eval_mod_sine_deg": self.sine_deg, "btp_eval_mod_double_angle": self.double_angle,
                "btp_eval_mod_arcsine_deg": 0, "btp_stc_start_level": self.stc_start, "btp_stc_depth": self.stc_depth,
                "btp_stc_bsgs_ratio": self.ratio, "btp_output_level": self.out_level}


def pipeline_rotations(spec):
    """rotation steps the product pipeline asks for (a subset of the frontend's rotations_for_bootstrapping())"""
    p = spec.params
    q = np.array(p.q, dtype=np.uint64)
    pp = np.array(p.p, dtype=np.uint64)
    rots = (C.c_int * 4096)()
    rc = lib().btp_oracle_run(spec.ints(), spec.doubles(), R.P(q), len(q), R.P(pp), len(pp), 0, None, None, 0, None, None, 0, None, None,
                              None, None, rots, 4096)
    assert rc == 0
    out = []
    for v in rots:
        if v < 0:
            break
        out.append(v)
    return out


def fixed_weight_secret(N, weight, seed):
    rng = np.random.default_rng(seed)
    s = np.zeros(N, dtype=np.int8)
    idx = rng.choice(N, size=weight, replace=False)
    s[idx] = rng.choice(np.array([-1, 1], dtype=np.int8), size=weight)
    return s


class BtpKeys:
    """dense secret (weight h), ephemeral sparse secret (weight 32), rlk, the Galois keys of the pipeline's rotations plus
    the row key, and the two encapsulation keys -- what CkksBtpContext generates on the reference side"""

    def __init__(self, orc, spec, seed=7, h_dense=192, h_sparse=32, rotations=None):
        self.o, self.spec = orc, spec
        N = orc.N
        top = len(orc.p.q) - 1
        self.top = top
        self.s = fixed_weight_secret(N, min(h_dense, N // 2), seed)
        self.s_sparse = fixed_weight_secret(N, h_sparse, seed + 1)
        self.rlk = orc.gen_rlk(self.s, seed + 2, top)
        rots = pipeline_rotations(spec) if rotations is None else rotations
        self.gals = [R.galois_for_col_step(N, r) for r in rots] + [R.galois_for_row(N)]
        self.glk = [orc.gen_glk(self.s, seed + 3 + g % 1000003, top, g) for g in self.gals]
        self.swk_dts = self._swk(self.s, self.s_sparse, seed + 4, 0)
        self.swk_std = self._swk(self.s_sparse, self.s, seed + 5, top)

    def _swk(self, s_in, s_out, seed, level):
        evk = np.zeros(self.o.evk_shape(level), dtype=np.uint64)
        lib().orc_gen_swk(self.o.hp, s_in.ctypes.data_as(i8p), s_out.ctypes.data_as(i8p), C.c_uint64(seed), level, R.P(evk))
        return evk


def encrypt_message(orc, keys, values, seed):
    """level-0 ciphertext of `values` (N/2 complex numbers) at the default scale"""
    pt = R.ckks_encode(orc, values, 0, orc.p.scale)
    return orc.encrypt(keys.s, seed, 0, pt)


def decode(orc, keys, ct, scale):
    return R.ckks_decode(orc, orc.decrypt(keys.s, ct), scale)


def oracle_bootstrap(orc, spec, keys, cts):
    """cts [B][2][1][N] -> ([B][2][out_level+1][N], stats) by the oracle run of the product pipeline"""
    p = orc.p
    q = np.array(p.q, dtype=np.uint64)
    pp = np.array(p.p, dtype=np.uint64)
    cts = np.ascontiguousarray(cts, dtype=np.uint64)
    B = cts.shape[0]
    out = np.zeros((B, 2, spec.out_level + 1, orc.N), dtype=np.uint64)
    gals = np.array(keys.gals, dtype=np.uint64)
    ptrs = (u64p * len(keys.glk))(*[R.P(k) for k in keys.glk])
    stats = (C.c_longlong * 8)()
    rc = lib().btp_oracle_run(spec.ints(), spec.doubles(), R.P(q), len(q), R.P(pp), len(pp), B, R.P(cts), R.P(keys.rlk), keys.top,
                              R.P(gals), ptrs, len(keys.glk), R.P(keys.swk_dts), R.P(keys.swk_std), R.P(out), stats, None, 0)
    if rc != 0:
        raise RuntimeError("btp_oracle_run failed")
    names = ["rotations", "key_switches", "mults", "rescales", "plain_mults"]
    return out, dict(zip(names, list(stats)[:5]))
