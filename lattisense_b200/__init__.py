"""lattisense_b200 — host-side Python mirror of the B200 mega_ag GPU executor.

The product is the C-ABI shared library ``_lib/libgpu_mega_ag_runner.so`` (hand-written sm_100a CUDA
behind ``include/lsb200.h`` and ``include/lattisense_gpu_runner.h``).  This package only binds it
with ctypes for tests, ``bench.py`` and ``__graft_entry__``; it contains no arithmetic and no CPU
fallback: if the library (or a CUDA device) is missing, calls fail loudly.

Reference interfaces mirrored: ``FheTaskGpu`` (cxx_sdk_v2/cxx_fhe_task_gpu.cpp:32-117) over
``create_/bind_/run_/release_fhe_gpu_task`` (mega_ag_runners/wrapper.h:67-85).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libgpu_mega_ag_runner.so")

SCHEME_BFV, SCHEME_CKKS = 0, 1

u64p = C.POINTER(C.c_uint64)
i32p = C.POINTER(C.c_int)

_lib = None


class LsbError(RuntimeError):
    pass


def lib():
    """Load the CUDA library.  Raises if it was not built — there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LsbError("%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(make -C lattisense_b200/csrc); there is no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.lsb_last_error.restype = C.c_char_p
    L.lsb_version.restype = C.c_char_p
    L.lsb_context_create.restype = C.c_void_p
    L.lsb_context_create.argtypes = [C.c_int, C.c_int, C.c_int, u64p, C.c_int, u64p, C.c_int, C.c_uint64]
    L.lsb_context_destroy.argtypes = [C.c_void_p]
    L.lsb_context_num_moduli.argtypes = [C.c_void_p]
    L.lsb_context_modulus.restype = C.c_uint64
    L.lsb_context_modulus.argtypes = [C.c_void_p, C.c_int]
    L.lsb_context_level_qmul.argtypes = [C.c_void_p, C.c_int]
    L.lsb_device_alloc.restype = C.c_void_p
    L.lsb_device_alloc.argtypes = [C.c_void_p, C.c_size_t]
    L.lsb_device_free.argtypes = [C.c_void_p, C.c_void_p]
    L.lsb_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    L.lsb_download.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    L.lsb_stream_sync.argtypes = [C.c_void_p, C.c_void_p]
    L.lsb_set_chunk.argtypes = [C.c_void_p, C.c_int]
    vp, ci = C.c_void_p, C.c_int
    L.lsb_ntt.argtypes = [vp, vp, vp, ci, i32p, ci, ci, ci, vp]
    L.lsb_mult.argtypes = [vp, ci, vp, vp, vp, ci, vp]
    L.lsb_relin.argtypes = [vp, ci, vp, vp, ci, vp, ci, vp]
    L.lsb_rescale.argtypes = [vp, ci, ci, vp, vp, ci, vp]
    L.lsb_drop_level.argtypes = [vp, ci, ci, vp, vp, ci, vp]
    L.lsb_mod_raise.argtypes = [vp, ci, ci, vp, vp, ci, vp]
    L.lsb_mod_raise_static.argtypes = [vp, ci, ci, vp, vp, ci, vp]
    L.lsb_linear_transform.argtypes = [vp, ci, vp, ci, vp, vp, ci, ci, vp, vp, ci, vp, vp]
    L.lsb_rotate.argtypes = [vp, ci, vp, vp, ci, C.c_uint64, vp, ci, vp]
    L.lsb_add.argtypes = [vp, ci, ci, vp, vp, vp, ci, vp]
    L.lsb_sub.argtypes = [vp, ci, ci, vp, vp, vp, ci, vp]
    L.lsb_neg.argtypes = [vp, ci, ci, vp, vp, ci, vp]
    L.lsb_add_plain.argtypes = [vp, ci, ci, vp, vp, vp, ci, ci, vp]
    L.lsb_add_plain_ringt.argtypes = [vp, ci, ci, vp, vp, vp, ci, ci, vp]
    L.lsb_mul_plain.argtypes = [vp, ci, ci, vp, vp, vp, ci, vp]
    L.lsb_mul_plain_ringt.argtypes = [vp, ci, ci, vp, vp, vp, ci, vp]
    L.lsb_mult_relin.argtypes = [vp, ci, vp, vp, vp, ci, vp, ci, vp]
    L.lsb_mult_relin_rescale.argtypes = [vp, ci, vp, vp, vp, ci, vp, ci, vp]
    L.lsb_ks_shard_layout.argtypes = [vp, ci, ci, ci, i32p, i32p, i32p, i32p, i32p, i32p]
    L.lsb_ks_shard_begin.argtypes = [vp, ci, ci, ci, vp, vp, ci, vp]
    L.lsb_ks_shard_inner.argtypes = [vp, ci, ci, ci, vp, vp, vp, vp, vp, ci, vp]
    L.lsb_ks_shard_begin_p2p.argtypes = [vp, ci, ci, ci, vp, vp, vp]
    L.lsb_ks_shard_inner_p2p.argtypes = [vp, ci, ci, ci, vp, vp, vp, vp, vp, vp]
    L.lsb_ks_shard_end.argtypes = [vp, ci, ci, ci, vp, vp, vp, vp, ci, vp]
    L.lsb_rotate_hoisted.argtypes = [vp, ci, vp, ci, C.POINTER(C.c_void_p), ci, u64p, C.POINTER(C.c_void_p), ci, vp]
    L.lsb_mac.argtypes = [vp, ci, ci, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), ci, vp, vp, ci, vp]
    L.lsb_bootstrap_create.restype = C.c_void_p
    L.lsb_bootstrap_create.argtypes = [vp, i32p, C.POINTER(C.c_double)]
    L.lsb_bootstrap_destroy.argtypes = [vp]
    L.lsb_bootstrap_rotations.argtypes = [vp, i32p, ci]
    L.lsb_bootstrap_run.argtypes = [vp, vp, vp, ci, u64p, C.POINTER(C.c_void_p), ci, vp, vp, vp, ci, vp, C.POINTER(C.c_longlong)]
    L.lsb_nccl_unique_id.argtypes = [vp]
    L.lsb_ks_shard_create.restype = C.c_void_p
    L.lsb_ks_shard_create.argtypes = [vp, ci, ci, ci, vp, ci]
    L.lsb_ks_shard_destroy.argtypes = [vp]
    L.lsb_ks_shard_run.argtypes = [vp, vp, vp, vp, vp, ci, vp]
    L.lsb_ks_shard_all_gather_bytes.restype = C.c_longlong
    L.lsb_ks_shard_all_gather_bytes.argtypes = [vp]
    L.lsb_ks_shard_last_error.restype = C.c_char_p
    L.lsb_launch_count.restype = C.c_longlong
    L.lsb_launch_count.argtypes = [vp]
    L.lsb_profile_enable.argtypes = [vp, ci]
    L.lsb_profile_read.argtypes = [vp, ci, C.POINTER(C.c_double), C.POINTER(C.c_longlong), C.POINTER(C.c_double)]
    _lib = L
    return L


STATIC_SCRATCH_WORDS = 1 << 19  # LSB_STATIC_SCRATCH_WORDS (include/lsb200.h)


def _check(rc):
    if rc != 0:
        raise LsbError(lib().lsb_last_error().decode() or "lsb call failed")


class DevBuf:
    """A device allocation of uint64 words owned by a Context."""

    def __init__(self, ctx, words):
        self.ctx, self.words = ctx, int(words)
        self.ptr = lib().lsb_device_alloc(ctx.h, self.words)
        if not self.ptr:
            raise LsbError(lib().lsb_last_error().decode())

    def free(self):
        if self.ptr:
            lib().lsb_device_free(self.ctx.h, self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Context:
    """RNS context on one GPU (init_gpu_context, mega_ag_runners/gpu/gpu_wrapper.cu:53-138)."""

    def __init__(self, scheme, log_n, q, p, t=0, device=0):
        self.scheme, self.logN, self.N = scheme, log_n, 1 << log_n
        self.q, self.p, self.t = list(q), list(p), t
        self.k = len(self.p)
        qa = np.array(self.q, dtype=np.uint64)
        pa = np.array(self.p, dtype=np.uint64)
        self.h = lib().lsb_context_create(device, scheme, log_n, qa.ctypes.data_as(u64p), len(qa),
                                          pa.ctypes.data_as(u64p), len(pa), t)
        if not self.h:
            raise LsbError(lib().lsb_last_error().decode())
        self.stream = None

    def close(self):
        if getattr(self, "h", None):
            lib().lsb_context_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- memory ----
    def alloc(self, words):
        return DevBuf(self, words)

    def upload(self, arr, buf=None):
        arr = np.ascontiguousarray(arr, dtype=np.uint64)
        if buf is None:
            buf = DevBuf(self, arr.size)
        _check(lib().lsb_upload(self.h, buf.ptr, arr.ctypes.data, arr.size, self.stream))
        _check(lib().lsb_stream_sync(self.h, self.stream))
        return buf

    def download(self, buf, shape):
        out = np.empty(shape, dtype=np.uint64)
        assert out.size <= buf.words
        _check(lib().lsb_download(self.h, out.ctypes.data, buf.ptr, out.size, self.stream))
        _check(lib().lsb_stream_sync(self.h, self.stream))
        return out

    def sync(self):
        _check(lib().lsb_stream_sync(self.h, self.stream))

    def set_chunk(self, chunk):
        lib().lsb_set_chunk(self.h, int(chunk))

    def launch_count(self):
        return int(lib().lsb_launch_count(self.h))

    def profile(self, on):
        lib().lsb_profile_enable(self.h, 1 if on else 0)

    def profile_read(self, kind):
        ms, n, b = C.c_double(), C.c_longlong(), C.c_double()
        _check(lib().lsb_profile_read(self.h, kind, C.byref(ms), C.byref(n), C.byref(b)))
        return ms.value, n.value, b.value

    def beta(self, level):
        return (level + 1 + self.k - 1) // self.k

    # ---- device ops (DevBuf in / DevBuf out, asynchronous on self.stream) ----
    def ntt_dev(self, d_in, d_out, rows, mod_idx, inverse, batch=1):
        mi = (C.c_int * len(mod_idx))(*mod_idx)
        _check(lib().lsb_ntt(self.h, d_in.ptr, d_out.ptr, rows, mi, len(mod_idx), 1 if inverse else 0, batch,
                             self.stream))

    def mult_dev(self, level, a, b, out3, batch=1):
        _check(lib().lsb_mult(self.h, level, a.ptr, b.ptr, out3.ptr, batch, self.stream))

    def relin_dev(self, level, ct3, rlk, key_level, out2, batch=1):
        _check(lib().lsb_relin(self.h, level, ct3.ptr, rlk.ptr, key_level, out2.ptr, batch, self.stream))

    def rescale_dev(self, level, npoly, a, out, batch=1):
        _check(lib().lsb_rescale(self.h, level, npoly, a.ptr, out.ptr, batch, self.stream))

    def drop_level_dev(self, level, npoly, a, out, batch=1):
        _check(lib().lsb_drop_level(self.h, level, npoly, a.ptr, out.ptr, batch, self.stream))

    def rotate_dev(self, level, ct, glk, key_level, gal, out, batch=1):
        _check(lib().lsb_rotate(self.h, level, ct.ptr, glk.ptr, key_level, gal, out.ptr, batch, self.stream))

    def mult_relin_dev(self, level, a, b, rlk, key_level, out, batch=1):
        _check(lib().lsb_mult_relin(self.h, level, a.ptr, b.ptr, rlk.ptr, key_level, out.ptr, batch, self.stream))

    def mult_relin_rescale_dev(self, level, a, b, rlk, key_level, out, batch=1):
        _check(lib().lsb_mult_relin_rescale(self.h, level, a.ptr, b.ptr, rlk.ptr, key_level, out.ptr, batch,
                                            self.stream))

    # ---- host convenience (numpy in / numpy out); arrays are [batch?][polys][limbs][N] ----
    def _run(self, fn, ins, out_shape):
        bufs = [self.upload(x) for x in ins]
        out = DevBuf(self, int(np.prod(out_shape)))
        fn(bufs, out)
        res = self.download(out, out_shape)
        for b in bufs:
            b.free()
        out.free()
        return res

    @staticmethod
    def _batch(a, nd):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        return (a[None], True) if a.ndim == nd else (a, False)

    def ntt(self, a, mod_idx, inverse=False):
        """a: [..., rows, N]; limb r uses modulus index mod_idx[r % len(mod_idx)]"""
        a = np.ascontiguousarray(a, dtype=np.uint64)
        rows = a.shape[-2]
        batch = int(np.prod(a.shape[:-2])) if a.ndim > 2 else 1
        return self._run(lambda b, o: self.ntt_dev(b[0], o, rows, list(mod_idx), inverse, batch), [a], a.shape)

    def mult(self, a, b):
        a, single = self._batch(a, 3)
        b, _ = self._batch(b, 3)
        B, _, l, N = a.shape
        r = self._run(lambda bf, o: self.mult_dev(l - 1, bf[0], bf[1], o, B), [a, b], (B, 3, l, N))
        return r[0] if single else r

    def relin(self, ct3, rlk):
        ct3, single = self._batch(ct3, 3)
        B, _, l, N = ct3.shape
        key_level = rlk.shape[2] - self.k - 1
        r = self._run(lambda bf, o: self.relin_dev(l - 1, bf[0], bf[1], key_level, o, B), [ct3, rlk], (B, 2, l, N))
        return r[0] if single else r

    def rescale(self, ct):
        ct, single = self._batch(ct, 3)
        B, npoly, l, N = ct.shape
        r = self._run(lambda bf, o: self.rescale_dev(l - 1, npoly, bf[0], o, B), [ct], (B, npoly, l - 1, N))
        return r[0] if single else r

    def mod_raise(self, ct0, top_level):
        """level-0 ciphertext [2][1][N] -> [2][top_level+1][N] (bootstrapping ModRaise)"""
        ct0, single = self._batch(ct0, 3)
        B, npoly, l, N = ct0.shape
        assert l == 1
        r = self._run(lambda bf, o: _check(lib().lsb_mod_raise(self.h, top_level, npoly, bf[0].ptr, o.ptr, B, self.stream)),
                      [ct0], (B, npoly, top_level + 1, N))
        return r[0] if single else r

    def mod_raise_static_dev(self, top_level, npoly, d_in, d_out, batch):
        """lsb_mod_raise_static on caller-owned device buffers: ModRaise that allocates nothing (scratch in the library's
        static device array of STATIC_SCRATCH_WORDS words; no counterpart in the reference)."""
        _check(lib().lsb_mod_raise_static(self.h, top_level, npoly, d_in.ptr, d_out.ptr, batch, self.stream))

    def linear_transform(self, ct, plan, n1, keys):
        """BSGS diagonal-form linear transform + rescale (lsb_linear_transform).  plan: [(diagonal index k, plaintext
        [l][N] NTT, slot values pre-rotated by -(k - k % n1))]; keys: {column rotation: evaluation key array}."""
        ct = np.ascontiguousarray(ct, dtype=np.uint64)
        _, l, N = ct.shape
        level = l - 1
        d_ct = self.upload(ct)
        d_pts = [self.upload(pt) for _, pt in plan]
        rots = sorted(keys)
        d_keys = [self.upload(keys[r]) for r in rots]
        key_level = (keys[rots[0]].shape[2] - self.k - 1) if rots else level
        n, nk = len(plan), len(rots)
        diag_rot = (C.c_int * n)(*[int(k) for k, _ in plan])
        diag_ptr = (C.c_void_p * n)(*[b.ptr for b in d_pts])
        key_rot = (C.c_int * max(nk, 1))(*rots)
        key_ptr = (C.c_void_p * max(nk, 1))(*[b.ptr for b in d_keys])
        out = DevBuf(self, 2 * level * N)
        _check(lib().lsb_linear_transform(self.h, level, d_ct.ptr, n, diag_rot, diag_ptr, n1, nk, key_rot, key_ptr, key_level,
                                          out.ptr, self.stream))
        res = self.download(out, (2, level, N))
        for b in [d_ct, out] + d_pts + d_keys:
            b.free()
        return res

    def drop_level(self, ct):
        ct, single = self._batch(ct, 3)
        B, npoly, l, N = ct.shape
        r = self._run(lambda bf, o: self.drop_level_dev(l - 1, npoly, bf[0], o, B), [ct], (B, npoly, l - 1, N))
        return r[0] if single else r

    def rotate(self, ct, glk, gal):
        ct, single = self._batch(ct, 3)
        B, _, l, N = ct.shape
        key_level = glk.shape[2] - self.k - 1
        r = self._run(lambda bf, o: self.rotate_dev(l - 1, bf[0], bf[1], key_level, gal, o, B), [ct, glk],
                      (B, 2, l, N))
        return r[0] if single else r

    def mult_relin(self, a, b, rlk, rescale=False):
        a, single = self._batch(a, 3)
        b, _ = self._batch(b, 3)
        B, _, l, N = a.shape
        key_level = rlk.shape[2] - self.k - 1
        if rescale:
            r = self._run(lambda bf, o: self.mult_relin_rescale_dev(l - 1, bf[0], bf[1], bf[2], key_level, o, B),
                          [a, b, rlk], (B, 2, l - 1, N))
        else:
            r = self._run(lambda bf, o: self.mult_relin_dev(l - 1, bf[0], bf[1], bf[2], key_level, o, B),
                          [a, b, rlk], (B, 2, l, N))
        return r[0] if single else r

    def _binary(self, name, a, b):
        a, single = self._batch(a, 3)
        b, _ = self._batch(b, 3)
        B, npoly, l, N = a.shape
        f = getattr(lib(), name)
        r = self._run(lambda bf, o: _check(f(self.h, l - 1, npoly, bf[0].ptr, bf[1].ptr, o.ptr, B, self.stream)),
                      [a, b], a.shape)
        return r[0] if single else r

    def add(self, a, b):
        return self._binary("lsb_add", a, b)

    def sub(self, a, b):
        return self._binary("lsb_sub", a, b)

    def neg(self, a):
        a, single = self._batch(a, 3)
        B, npoly, l, N = a.shape
        r = self._run(lambda bf, o: _check(lib().lsb_neg(self.h, l - 1, npoly, bf[0].ptr, o.ptr, B, self.stream)),
                      [a], a.shape)
        return r[0] if single else r

    def _plain(self, name, ct, pt, *extra):
        ct, single = self._batch(ct, 3)
        pt = np.ascontiguousarray(pt, dtype=np.uint64)
        if pt.ndim == 2:
            pt = pt[None]
        B, npoly, l, N = ct.shape
        f = getattr(lib(), name)
        r = self._run(lambda bf, o: _check(f(self.h, l - 1, npoly, bf[0].ptr, bf[1].ptr, o.ptr, *extra, B,
                                             self.stream)), [ct, pt], ct.shape)
        return r[0] if single else r

    def add_plain(self, ct, pt, subtract=False):
        return self._plain("lsb_add_plain", ct, pt, 1 if subtract else 0)

    def add_plain_ringt(self, ct, pt, subtract=False):
        return self._plain("lsb_add_plain_ringt", ct, pt, 1 if subtract else 0)

    def mul_plain(self, ct, pt):
        return self._plain("lsb_mul_plain", ct, pt)

    def mul_plain_ringt(self, ct, pt):
        return self._plain("lsb_mul_plain_ringt", ct, pt)

    def rotate_hoisted(self, ct, glks, gals):
        """[rotate(ct, glk, gal) for glk, gal in zip(glks, gals)] with one shared decomposition (lsb_rotate_hoisted)"""
        ct = np.ascontiguousarray(ct, dtype=np.uint64)
        batched = ct.ndim == 4
        B = ct.shape[0] if batched else 1
        level = ct.shape[-2] - 1
        key_level = glks[0].shape[2] - self.k - 1
        d = self.upload(ct)
        dk = [self.upload(k) for k in glks]
        outs = [DevBuf(self, ct.size) for _ in glks]
        pk = (C.c_void_p * len(dk))(*[x.ptr for x in dk])
        po = (C.c_void_p * len(outs))(*[x.ptr for x in outs])
        ga = np.array(gals, dtype=np.uint64)
        _check(lib().lsb_rotate_hoisted(self.h, level, d.ptr, len(dk), pk, key_level, ga.ctypes.data_as(u64p), po, B, self.stream))
        return [self.download(o, ct.shape) for o in outs]

    def mac(self, cts, pts, ringt=False, acc=None):
        """sum_i cts[i] x pts[i] (+ acc): bind_gpu_cmp_sum / bind_gpu_cmpac_sum.  cts[i] [2][l][N] (or [B][2][l][N])."""
        cts = [np.ascontiguousarray(c, dtype=np.uint64) for c in cts]
        batched = cts[0].ndim == 4
        B = cts[0].shape[0] if batched else 1
        level = cts[0].shape[-2] - 1
        dc = [self.upload(c) for c in cts]
        dp = [self.upload(p) for p in pts]
        da = self.upload(acc) if acc is not None else None
        out = DevBuf(self, cts[0].size)
        pc = (C.c_void_p * len(dc))(*[d.ptr for d in dc])
        pp = (C.c_void_p * len(dp))(*[d.ptr for d in dp])
        _check(lib().lsb_mac(self.h, level, len(dc), pc, pp, 1 if ringt else 0, da.ptr if da else None, out.ptr, B, self.stream))
        return self.download(out, cts[0].shape)


class Bootstrapper:
    """CKKS bootstrapping on the device (the `bootstrap` node, mega_ag_executors_gpu.cu:410-426) through lsb_bootstrap_*.
    iparams / dparams as in include/lsb200.h.  Keys are uploaded once with set_keys and stay on the device."""

    def __init__(self, ctx, iparams, dparams):
        self.ctx = ctx
        self.ip = (C.c_int * 11)(*iparams)
        self.dp = (C.c_double * 4)(*dparams)
        self.out_level = int(iparams[10])
        self.h = lib().lsb_bootstrap_create(ctx.h, self.ip, self.dp)
        if not self.h:
            raise LsbError(lib().lsb_last_error().decode())
        self.keys = None

    def rotations(self):
        buf = (C.c_int * 4096)()
        n = lib().lsb_bootstrap_rotations(self.h, buf, 4096)
        if n < 0:
            raise LsbError(lib().lsb_last_error().decode())
        return list(buf[:n])

    def set_keys(self, rlk, gals, glks, swk_dts, swk_std):
        c = self.ctx
        self.keys = {"rlk": c.upload(rlk), "glk": [c.upload(k) for k in glks], "dts": c.upload(swk_dts), "std": c.upload(swk_std),
                     "gals": np.array(gals, dtype=np.uint64), "key_level": rlk.shape[2] - c.k - 1}

    def run(self, cts):
        """cts [B][2][1][N] uint64 -> ([B][2][out_level+1][N], stats)"""
        c, k = self.ctx, self.keys
        cts = np.ascontiguousarray(cts, dtype=np.uint64)
        B = cts.shape[0]
        d_in = c.upload(cts)
        shape = (B, 2, self.out_level + 1, c.N)
        d_out = DevBuf(c, int(np.prod(shape)))
        ptrs = (C.c_void_p * len(k["glk"]))(*[g.ptr for g in k["glk"]])
        stats = (C.c_longlong * 5)()
        _check(lib().lsb_bootstrap_run(self.h, d_in.ptr, k["rlk"].ptr, k["key_level"], k["gals"].ctypes.data_as(u64p), ptrs,
                                       len(k["glk"]), k["dts"].ptr, k["std"].ptr, d_out.ptr, B, c.stream, stats))
        out = c.download(d_out, shape)
        return out, dict(zip(["rotations", "key_switches", "mults", "rescales", "plain_mults"], list(stats)))

    def close(self):
        if getattr(self, "h", None):
            lib().lsb_bootstrap_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
