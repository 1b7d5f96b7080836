"""Host-side tests of the mega_ag runner boundary (no GPU): exported symbols, graph loading with ABI-bridge
insertion (mega_ag_runners/mega_ag.cpp:388-544), chain fusion, component sharding, error reporting."""
import ctypes as C
import json
import os
import re

import pytest

import graph_eval as GE

ROOT = GE.ROOT


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"//[^\n]*", "", txt)
    txt = re.sub(r"typedef[^;{]*\([^;]*;", "", txt)  # function-pointer typedefs
    names = re.findall(r"\b([a-z_][a-z0-9_]*)\s*\([^;{]*\)\s*;", txt)
    return sorted({n for n in names if n not in ("defined", "sizeof") and not n.endswith("_t")})


@pytest.mark.parametrize("header", ["lsb200.h", "lattisense_gpu_runner.h"])
def test_library_exports_every_declared_symbol(header):
    lib = C.CDLL(os.path.join(GE.LIB_DIR, "libgpu_mega_ag_runner.so"))
    names = _declared(header)
    assert len(names) >= 9
    for n in names:
        assert hasattr(lib, n), "missing export " + n
    if header == "lattisense_gpu_runner.h":
        for n in ("create_fhe_gpu_task", "release_fhe_gpu_task", "bind_gpu_task_abi_bridge_executors",
                  "bind_gpu_task_custom_executors", "run_fhe_gpu_task"):
            assert n in names


def _describe(case):
    h = GE.harness()
    t = h.th_open(os.path.join(GE.GRAPHS, case).encode(), 1)
    assert t, h.th_last_error().decode()
    out = (C.c_longlong * 8)()
    assert GE.runner_lib().lsb_gpu_task_describe(t, out) == 0
    return h, t, dict(zip(["computes", "data", "host", "load", "store", "gpu", "fused", "components"], list(out)))


def test_bridge_nodes_cfg1(monkeypatch):
    # cfg1: 2 compute nodes, inputs x, y, rlk, one output:
    # + 3 EXPORT + 3 LOAD + 1 STORE + 1 IMPORT = 10 nodes; mult+relin fuse into one GPU unit
    monkeypatch.setenv("LSB_NO_HOST_FUSE", "1")
    h, t, d = _describe("cfg1_bfv_mult")
    try:
        assert d["computes"] == 10
        assert (d["host"], d["load"], d["store"], d["gpu"], d["fused"]) == (4, 3, 1, 1, 1)
    finally:
        h.th_close(t)
    # default plan: each EXPORT runs inside its LOAD unit and the IMPORT inside the STORE unit (host chains)
    monkeypatch.delenv("LSB_NO_HOST_FUSE")
    h, t, d = _describe("cfg1_bfv_mult")
    try:
        assert d["computes"] == 10
        assert (d["host"], d["load"], d["store"], d["gpu"], d["fused"]) == (0, 3, 1, 1, 1)
        assert d["data"] == 5 + 3 * 2 + 2
        assert d["components"] == 1
    finally:
        h.th_close(t)


def test_components_and_sharding_cfg2():
    h, t, d = _describe("cfg2_ckks_mult_x8")
    try:
        # 8 independent mult->relin->rescale chains sharing only the rlk
        assert d["components"] == 8 and d["gpu"] == 8 and d["fused"] == 16
        assert d["load"] == 17 and d["store"] == 8
        slots = (C.c_int * 8)()
        for n in (1, 2, 4, 8):
            assert GE.runner_lib().lsb_gpu_task_shard_preview(t, n, slots, 8) == 8
            counts = [list(slots).count(s) for s in range(n)]
            assert counts == [8 // n] * n, counts
    finally:
        h.th_close(t)


def test_small_graphs_open_fewer_devices(monkeypatch, tmp_path):
    """lsb_gpu_task_set_devices is a request: 8 small components stay on one device, cfg3 x 64 uses 4 of 8 (the measured
    optimum), and the cap can be switched off"""
    lib = GE.runner_lib()
    lib.lsb_gpu_task_worthwhile_devices.argtypes = [C.c_void_p, C.c_int]
    monkeypatch.delenv("LSB_MIN_WEIGHT_PER_DEVICE", raising=False)
    h = GE.harness()
    for case, want in (("cfg2_ckks_mult_x8", {1: 1, 2: 1, 8: 1}), ("cfg3_bfv_poly7_x64", {1: 1, 2: 2, 4: 4, 8: 4})):
        path = os.path.join(GE.GRAPHS, case)
        if not os.path.exists(os.path.join(path, "mega_ag.json")):  # large fixtures are committed gzipped
            path = GE.unpack_case(case, str(tmp_path))
        t = h.th_open(path.encode(), 1)
        assert t, h.th_last_error().decode()
        try:
            for n, w in want.items():
                assert lib.lsb_gpu_task_worthwhile_devices(t, n) == w, (case, n)
            monkeypatch.setenv("LSB_MIN_WEIGHT_PER_DEVICE", "0")
            assert lib.lsb_gpu_task_worthwhile_devices(t, 8) == 8
            monkeypatch.delenv("LSB_MIN_WEIGHT_PER_DEVICE")
        finally:
            h.th_close(t)


def test_custom_node_splits_device_path():
    h, t, d = _describe("bfv8k_custom_in_middle")
    try:
        # x,y,rlk loaded; (mult,relin) fused; result stored + imported for the custom node; custom output exported
        # + loaded again for the add; final output stored + imported
        assert d["gpu"] == 2 and d["fused"] == 1
        assert d["store"] == 2 and d["load"] == 4
        assert d["host"] == 1  # the custom node; the 4 exports / 2 imports run inside their LOAD / STORE units
    finally:
        h.th_close(t)


@pytest.mark.parametrize("case", GE.all_cases())
def test_every_fixture_graph_loads(case):
    h, t, d = _describe(case)
    try:
        g = json.load(open(os.path.join(GE.GRAPHS, case, "mega_ag.json")))
        n_in, n_out = len(g["inputs"]), len(g["outputs"])
        custom = sum(1 for c in g["compute"].values() if c.get("is_custom"))
        assert d["computes"] >= len(g["compute"]) + 2 * n_in + 2 * n_out
        assert d["gpu"] + d["fused"] == len(g["compute"]) - custom
        assert d["store"] >= n_out and d["load"] >= n_in
    finally:
        h.th_close(t)


def test_missing_task_directory_is_an_error():
    h = GE.harness()
    assert not h.th_open(b"/nonexistent/task", 1)
    assert "Cannot open MegaAG file" in h.th_last_error().decode()


def test_run_without_gpu_fails_loudly():
    """No CPU fallback: on a box without CUDA the run must fail, not compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    case = GE.GraphCase(os.path.join(GE.GRAPHS, "ckks8k_cac"))
    with pytest.raises(RuntimeError, match="no CUDA device|CUDA|cuda"):
        case.run()


def test_sum_reassociation_in_the_plan(tmp_path, monkeypatch):
    # packed convolution, 64x64 input / 5x5 kernel: 50 ct x pt products accumulated through 49 dependent additions
    # (+ rescale + bias): the products and the additions collapse into ONE n-ary multiply-accumulate unit
    h, t, d = _describe("conv_64x64_5x5")
    try:
        assert d["gpu"] == 49 + 1 + 1 + 1          # rotations, the sum, rescale, + bias
        assert d["fused"] == 50 + 49 - 1
    finally:
        h.th_close(t)
    monkeypatch.setenv("LSB_NO_FUSE", "1")
    h, t, d = _describe("conv_64x64_5x5")
    try:
        assert (d["gpu"], d["fused"]) == (150, 0)
    finally:
        h.th_close(t)
    monkeypatch.delenv("LSB_NO_FUSE")
    # BFV degree-7 polynomial: products of ct x pt_mul are not folded (they need their own NTT round trip), the six
    # additions between them are one sum
    h, t, d = _describe("cfg3_bfv_poly7")
    try:
        assert d["fused"] >= 5
    finally:
        h.th_close(t)
    # the 38 400-node stress graph: 12 799 rotations + 1 sum + rescale + bias
    h = GE.harness()
    t = h.th_open(GE.unpack_case("conv_4x4_5x5_stress", str(tmp_path)).encode(), 1)
    assert t, h.th_last_error().decode()
    try:
        out = (C.c_longlong * 8)()
        assert GE.runner_lib().lsb_gpu_task_describe(t, out) == 0
        assert out[5] == 12799 + 3 and out[6] == 12800 + 12799 - 1
    finally:
        h.th_close(t)
