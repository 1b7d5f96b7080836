"""Generates tests/golden/graphs/<case>/{mega_ag.json,task_signature.json} with the REFERENCE's own
graph compiler (frontend/custom_task.py, imported from /root/reference — it runs in the build
container but does not travel to the GPU box, hence the committed fixtures).

    PYTHONPATH=/root/reference python tests/golden/make_graphs.py

Cases mirror the reference's GPU test list (unittests/test_gpu_bfv.py, test_gpu_ckks.py: N_OP
independent copies per graph) and the BASELINE.json configs.  Parameter sets: the reference's test
fixtures (unittests/fixture.hpp:74-79, 96-103) at n=8192 and the defaults at n=16384.
"""
import os
import shutil
import sys

REF = os.environ.get("LATTISENSE_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
from frontend.custom_task import *  # noqa: E402,F401,F403

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "graphs")
N_OP = 4

BFV8K = dict(n=8192, q=[0x3FFFFFFFEF8001, 0x4000000011C001, 0x40000000120001], p=[0x7FFFFFFFFB4001], t=65537)
CKKS8K = dict(n=8192, q=[0x1FFFEC001, 0x3FFF4001, 0x3FFE8001, 0x40020001, 0x40038001, 0x3FFC0001], p=[0x800004001],
              scale=float(1 << 30))


def bfv8k():
    return BfvParam.create_custom_param(**BFV8K)


def ckks8k():
    return CkksParam.create_custom_param(n=CKKS8K["n"], q=list(CKKS8K["q"]), p=list(CKKS8K["p"]), scale=CKKS8K["scale"])


def emit(case, ins, outs, offline=()):
    d = os.path.join(OUT, case)
    shutil.rmtree(d, ignore_errors=True)
    process_custom_task(input_args=ins, offline_input_args=list(offline), output_args=outs,
                        output_instruction_path=d, fpga_acc=False)
    for f in os.listdir(d):
        if f not in ("mega_ag.json", "task_signature.json"):
            p = os.path.join(d, f)
            shutil.rmtree(p) if os.path.isdir(p) else os.remove(p)
    print("wrote", case)


def binary_case(case, param, Ct, Other, fn, lv, other_level=True):
    set_fhe_param(param)
    xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
    ys = [Other(f"y_{i}", level=lv) if other_level else Other(f"y_{i}") for i in range(N_OP)]
    zs = [fn(xs[i], ys[i], f"z_{i}") for i in range(N_OP)]
    emit(case, [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", zs)])


def gen_scheme(tag, mkparam, Ct, Pt, PtRingt, PtMul, lv, n_sp):
    binary_case(f"{tag}_cap", mkparam(), Ct, Pt, add, lv)
    binary_case(f"{tag}_cac", mkparam(), Ct, Ct, add, lv)
    binary_case(f"{tag}_csp", mkparam(), Ct, Pt, sub, lv)
    binary_case(f"{tag}_csc", mkparam(), Ct, Ct, sub, lv)
    binary_case(f"{tag}_cmp_ringt", mkparam(), Ct, PtRingt, mult, lv, other_level=False)
    binary_case(f"{tag}_cmp", mkparam(), Ct, PtMul, mult, lv)
    binary_case(f"{tag}_cmc", mkparam(), Ct, Ct, mult, lv)
    binary_case(f"{tag}_cmc_relin", mkparam(), Ct, Ct, mult_relin, lv)
    if tag.startswith("bfv"):
        binary_case(f"{tag}_cap_ringt", mkparam(), Ct, PtRingt, add, lv, other_level=False)
        binary_case(f"{tag}_csp_ringt", mkparam(), Ct, PtRingt, sub, lv, other_level=False)

    set_fhe_param(mkparam())
    xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
    zs = [neg(xs[i], f"z_{i}") for i in range(N_OP)]
    emit(f"{tag}_cneg", [Argument("in_x_list", xs)], [Argument("out_z_list", zs)])

    set_fhe_param(mkparam())
    xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
    ys = [Ct(f"y_{i}", level=lv) for i in range(N_OP)]
    zs = [rescale(mult_relin(xs[i], ys[i]), f"z_{i}") for i in range(N_OP)]
    emit(f"{tag}_cmc_relin_rescale", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", zs)])

    set_fhe_param(mkparam())
    xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
    zs = [rescale(mult_relin(xs[i], xs[i]), f"z_{i}") for i in range(N_OP)]
    emit(f"{tag}_csqr_relin_rescale", [Argument("in_x_list", xs)], [Argument("out_z_list", zs)])

    set_fhe_param(mkparam())
    xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
    zs = [rescale(xs[i], f"z_{i}") for i in range(N_OP)]
    emit(f"{tag}_rescale", [Argument("in_x_list", xs)], [Argument("out_z_list", zs)])

    set_fhe_param(mkparam())
    xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
    steps = [1, 2, 3, 5]
    ys = [rotate_cols(xs[i], steps) for i in range(N_OP)]
    emit(f"{tag}_rotate_col", [Argument("in_x_list", xs)], [Argument("out_y_list", ys)])

    set_fhe_param(mkparam())
    xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
    steps = [-900, 20, 400, 2000, 3009] if tag.startswith("bfv") else [-500, 20, 200, 2000, 4000]
    ys = [advanced_rotate_cols(xs[i], steps) for i in range(N_OP)]
    emit(f"{tag}_advanced_rotate_col", [Argument("in_x_list", xs)], [Argument("out_y_list", ys)])

    set_fhe_param(mkparam())
    xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
    ys = [rotate_rows(xs[i], f"y_{i}") for i in range(N_OP)]
    emit(f"{tag}_rotate_row", [Argument("in_x_list", xs)], [Argument("out_y_list", ys)])

    # ct x pt_ringt multiply-accumulate (cmp_sum / cmpac_sum nodes)
    set_fhe_param(mkparam())
    m = 20
    xs = [Ct(f"x_{i}", level=lv) for i in range(m)]
    ys = [PtRingt(f"y_{i}") for i in range(m)]
    z = ct_pt_mult_accumulate(xs, ys)
    emit(f"{tag}_ct_pt_ringt_mac", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z", z)])

    # ((x*y relin) rotated by 1) + x : mult -> relin -> rotate -> add chain
    set_fhe_param(mkparam())
    xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
    ys = [Ct(f"y_{i}", level=lv) for i in range(N_OP)]
    zs = []
    for i in range(N_OP):
        w = mult_relin(xs[i], ys[i])
        r = rotate_cols(w, 1)[0]
        zs.append(add(r, xs[i], f"z_{i}"))
    emit(f"{tag}_ctc_rotate_cac", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", zs)])

    # a custom (CPU) node in the middle of the graph: x*y -> custom -> + x
    set_fhe_param(mkparam())
    x, y = Ct("x", level=lv), Ct("y", level=lv)
    w = mult_relin(x, y)
    c = Ct("c", level=lv)
    custom_compute([w], c, "passthrough", {"note": "identity on the host", "k": 3})
    z = add(c, x, "z")
    emit(f"{tag}_custom_in_middle", [Argument("x", x), Argument("y", y)], [Argument("z", z)])


def gen_ckks_extra(tag, mkparam, lv):
    set_fhe_param(mkparam())
    xs = [CkksCiphertextNode(f"x_{i}", level=lv) for i in range(N_OP)]
    zs = [drop_level(xs[i], 2, f"z_{i}") for i in range(N_OP)]
    emit(f"{tag}_drop_level", [Argument("in_x_list", xs)], [Argument("out_z_list", zs)])


def gen_configs():
    # cfg1: BFV mult_relin n=16384 t=65537 level 2 (examples/bfv_mult_cpu)
    set_fhe_param(BfvParam.create_default_param(n=16384))
    x, y = BfvCiphertextNode("x", level=2), BfvCiphertextNode("y", level=2)
    z = mult_relin(x, y, "z")
    emit("cfg1_bfv_mult", [Argument("x", x), Argument("y", y)], [Argument("z", z)])

    # cfg2 (reduced copy count; bench.py replicates the node pattern to 1024): CKKS mult_relin + rescale
    set_fhe_param(CkksParam.create_default_param(n=16384))
    xs = [CkksCiphertextNode(f"x_{i}", level=3) for i in range(8)]
    ys = [CkksCiphertextNode(f"y_{i}", level=3) for i in range(8)]
    zs = [rescale(mult_relin(xs[i], ys[i]), f"z_{i}") for i in range(8)]
    emit("cfg2_ckks_mult_x8", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", zs)])

    # cfg3: BFV degree-7 polynomial (examples/bfv_poly_7_cpu/bfv_poly_7_cpu.py:31-53)
    set_fhe_param(BfvParam.create_default_param(n=16384))
    level = 4
    x = BfvCiphertextNode("x", level)
    a = [BfvPlaintextMulNode(f"a_{i}", level - 3) for i in range(1, 8)]
    a0 = BfvPlaintextNode("a_0", level - 3)
    x2 = rescale(mult_relin(x, x))
    x3 = rescale(mult_relin(x2, rescale(x)))
    x4 = rescale(mult_relin(x2, x2))
    x2r = rescale(x2)
    x5 = rescale(mult_relin(x3, x2r))
    x6 = rescale(mult_relin(x3, x3))
    x7 = rescale(mult_relin(x3, x4))
    pw = [rescale(rescale(rescale(x))), rescale(rescale(x2)), rescale(x3), rescale(x4), x5, x6, x7]
    acc = None
    for i in range(7):
        term = mult(pw[i], a[i])
        acc = term if acc is None else add(acc, term)
    yv = add(acc, a0, "y")
    emit("cfg3_bfv_poly7", [Argument("x", x), Argument("a", a), Argument("a_0", a0)], [Argument("y", yv)])

    # cfg4: CKKS logistic-regression inference (examples/ckks_logistic_regression_cpu/...py:26-59), n=16384 as shipped
    set_fhe_param(CkksParam.create_default_param(n=16384))
    level = 3
    xe = CkksCiphertextNode("x", level)
    w = CkksPlaintextRingtNode("w")
    b = CkksPlaintextNode("b", level - 1)
    s = rescale(mult(xe, w))
    for st in (16, 8, 4, 2, 1):
        s = add(s, rotate_cols(s, st)[0])
    s = add(s, b)
    half = CkksPlaintextRingtNode("half")
    o = rescale(mult(s, half), "out")
    emit("cfg4_ckks_logreg", [Argument("x", xe), Argument("w", w), Argument("b", b), Argument("half", half)],
         [Argument("out", o)])


def two_sided(x, reach, unit):
    """x rotated by -reach*unit .. +reach*unit (x itself in the middle)."""
    steps = [-i * unit for i in range(1, reach + 1)] + [i * unit for i in range(1, reach + 1)]
    r = rotate_cols(x, steps) if steps else []
    return list(reversed(r[:reach])) + [x] + r[reach:]


def gen_conv(case, shape, kernel, level=2):
    """Single-channel packed 2-D convolution as the reference's rotation benchmark builds it
    (examples/benchmark_convolution/benchmark_convolution.py:44-161): the input ciphertext holds
    pack = slots/(h*w) channels; every channel offset and every kernel tap is one rotate_cols, every
    tap a ct x pt product, all summed, rescaled, bias added.  (4,4)x(5,5) is SURVEY 8f-3's 38 400-node
    scheduler / rotation stress graph."""
    set_fhe_param(CkksParam.create_default_param(n=16384))
    h, w = shape
    kh, kw = kernel
    pack = (16384 // 2) // (h * w)
    x = CkksCiphertextNode("input_0_0", level=level)
    wts = [[CkksPlaintextNode(f"w_{c}_{k}", level) for k in range(kh * kw)] for c in range(pack)]
    bias = CkksPlaintextNode("b_0", level - 1)
    chans = [x] + (rotate_cols(x, [i * h * w for i in range(1, pack)]) if pack > 1 else [])
    acc = None
    for c, xc in enumerate(chans):
        taps = []
        for r in two_sided(xc, kw // 2, w):
            taps += two_sided(r, kh // 2, 1)
        for k, tap in enumerate(taps):
            prod = mult(tap, wts[c][k])
            acc = prod if acc is None else add(acc, prod)
    y = add(rescale(acc), bias, "y")
    emit(case, [Argument("input_0", [x]), Argument("w", wts), Argument("b", [bias])], [Argument("output", [y])])


def gen_logreg_n65536():
    """cfg4 at the ring degree BASELINE.json quotes it on (N=2^16): same node pattern as cfg4_ckks_logreg, moduli of
    CkksBtpParam.create_default_param() (frontend/custom_task.py:387-420; 25 Q + 5 P limbs, the graph uses level 3)."""
    set_fhe_param(CkksBtpParam.create_default_param())
    level = 3
    xe = CkksCiphertextNode("x", level)
    w = CkksPlaintextRingtNode("w")
    b = CkksPlaintextNode("b", level - 1)
    s = rescale(mult(xe, w))
    for st in (16, 8, 4, 2, 1):
        s = add(s, rotate_cols(s, st)[0])
    s = add(s, b)
    half = CkksPlaintextRingtNode("half")
    o = rescale(mult(s, half), "out")
    emit("cfg4_ckks_logreg_n65536", [Argument("x", xe), Argument("w", w), Argument("b", b), Argument("half", half)],
         [Argument("out", o)])


def gen_seal_flavoured():
    """Rotations as the SEAL plug-in's frontend emits them (custom_task.py:1586-1740): Galois generator 3 instead of 5, the
    element carried by the key datum - the runner must take it from there, not from the step."""
    for tag, mkparam, Ct, lv in (("bfv8k", bfv8k, BfvCiphertextNode, 2), ("ckks8k", ckks8k, CkksCiphertextNode, 4)):
        set_fhe_param(mkparam())
        xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
        ys = [seal_rotate_cols(xs[i], [1, 6, -3]) for i in range(N_OP)]
        zs = [seal_rotate_rows(xs[i], f"z_{i}") for i in range(N_OP)]
        ws = [seal_advanced_rotate_cols(xs[i], [20, -500]) for i in range(N_OP)]
        emit(f"{tag}_seal_rotate", [Argument("in_x_list", xs)],
             [Argument("out_y_list", ys), Argument("out_z_list", zs), Argument("out_w_list", ws)])


def gen_cfg3_replicas(copies=8):
    """cfg3 as BASELINE.json shards it over 2/4/8 GPUs: `copies` independent degree-7 polynomial evaluations (one DAG
    component each; a single instance has a critical path of 3 mult_relin and cannot use several GPUs), shared rlk."""
    set_fhe_param(BfvParam.create_default_param(n=16384))
    level = 4
    a = [BfvPlaintextMulNode(f"a_{i}", level - 3) for i in range(1, 8)]
    a0 = BfvPlaintextNode("a_0", level - 3)
    xs, ys = [], []
    for b in range(copies):
        x = BfvCiphertextNode(f"x_{b}", level)
        x2 = rescale(mult_relin(x, x))
        x3 = rescale(mult_relin(x2, rescale(x)))
        x4 = rescale(mult_relin(x2, x2))
        x5 = rescale(mult_relin(x3, rescale(x2)))
        x6 = rescale(mult_relin(x3, x3))
        x7 = rescale(mult_relin(x3, x4))
        pw = [rescale(rescale(rescale(x))), rescale(rescale(x2)), rescale(x3), rescale(x4), x5, x6, x7]
        acc = None
        for i in range(7):
            term = mult(pw[i], a[i])
            acc = term if acc is None else add(acc, term)
        xs.append(x)
        ys.append(add(acc, a0, f"y_{b}"))
    emit(f"cfg3_bfv_poly7_x{copies}", [Argument("x", xs), Argument("a", a), Argument("a_0", a0)], [Argument("y", ys)])


def gen_bootstrap():
    """The reference's bootstrap GPU cases (unittests/test_gpu_ckks.py:566-614 / test_gpu_ckks.cpp:743-808): toy_bootstrap
    (n = 8192), bootstrap and cmc_relin_rescale_bootstrap (n = 65536), on the N16QP1546H192H32 chain."""
    set_fhe_param(CkksBtpParam.create_toy_param())
    xs = [CkksCiphertextNode(f"x_{i}", level=0) for i in range(2)]
    ys = [bootstrap(xs[i], f"y_{i}") for i in range(2)]
    emit("ckks_btp8k_toy_bootstrap", [Argument("in_x_list", xs)], [Argument("out_y_list", ys)])

    set_fhe_param(CkksBtpParam.create_default_param())
    xs = [CkksCiphertextNode("x_0", level=0)]
    ys = [bootstrap(xs[0], "y_0")]
    emit("ckks_btp64k_bootstrap", [Argument("in_x_list", xs)], [Argument("out_y_list", ys)])

    p = CkksBtpParam.create_toy_param()
    set_fhe_param(p)
    lv = 3
    xs = [CkksCiphertextNode("x_0", level=lv)]
    ys = [CkksCiphertextNode("y_0", level=lv)]
    z = rescale(mult_relin(xs[0], ys[0]))
    zd = drop_level(z, z.level)
    res = [bootstrap(zd, "result_0")]
    emit("ckks_btp8k_cmc_relin_rescale_bootstrap", [Argument("in_x_list", xs), Argument("in_y_list", ys)],
         [Argument("out_z_list", res)])


def gen_reference_extras():
    """Reference GPU cases that had no fixture: unittests/test_gpu_bfv.py:116-549, test_gpu_ckks.py:233."""
    for tag, mkparam, Ct, lv in (("bfv8k", bfv8k, BfvCiphertextNode, 2), ("ckks8k", ckks8k, CkksCiphertextNode, 4)):
        set_fhe_param(mkparam())
        xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
        emit(f"{tag}_casc", [Argument("in_x_list", xs)], [Argument("out_z_list", [add(xs[i], xs[i], f"z_{i}") for i in range(N_OP)])])
        set_fhe_param(mkparam())
        xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
        emit(f"{tag}_cssc", [Argument("in_x_list", xs)], [Argument("out_z_list", [sub(xs[i], xs[i], f"z_{i}") for i in range(N_OP)])])
        set_fhe_param(mkparam())
        xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
        emit(f"{tag}_csqr", [Argument("in_x_list", xs)], [Argument("out_z_list", [mult(xs[i], xs[i], f"z_{i}") for i in range(N_OP)])])
        set_fhe_param(mkparam())
        xs = [Ct(f"x_{i}", level=lv) for i in range(N_OP)]
        emit(f"{tag}_csqr_relin", [Argument("in_x_list", xs)],
             [Argument("out_z_list", [mult_relin(xs[i], xs[i], f"z_{i}") for i in range(N_OP)])])

    bfv16 = lambda: BfvParam.create_default_param(n=16384)
    Ct, lv = BfvCiphertextNode, 3
    set_fhe_param(bfv16())
    xs = [Ct(f"x_{i}", level=lv) for i in range(4)]
    ys = [Ct(f"y_{i}", level=lv) for i in range(4)]
    tmp0 = mult_relin(xs[0], ys[0], "tmp_0")
    zs = [tmp0, mult_relin(tmp0, xs[1], "z_0"), mult_relin(xs[1], ys[1], "z_1"), mult_relin(xs[2], ys[2], "z_2"),
          mult_relin(xs[3], ys[3], "z_3")]
    emit("bfv16k_ctc_ctc_0", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", zs)])

    set_fhe_param(bfv16())
    xs = [Ct(f"x_{i}", level=lv) for i in range(4)]
    ys = [Ct(f"y_{i}", level=lv) for i in range(4)]
    t = [mult_relin(xs[i], ys[i], f"t_{i}") for i in range(4)]
    zs = [mult_relin(t[0], t[1], "z_0"), mult_relin(t[1], xs[2], "z_1"), mult_relin(t[2], xs[3], "z_2"), mult_relin(t[2], t[3], "z_3")]
    emit("bfv16k_ctc_ctc_1", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", zs)])

    set_fhe_param(bfv16())
    xs = [Ct("x_0", level=lv)]
    t = mult_relin(xs[0], xs[0], "x^2_0")
    emit("bfv16k_1_square_square", [Argument("in_x_list", xs)], [Argument("out_z_list", [mult_relin(t, t, "x^4_0")])])

    set_fhe_param(bfv16())
    xs = [Ct(f"x_{i}", level=lv) for i in range(3)]
    emit("bfv16k_double", [Argument("in_x_list", xs)],
         [Argument("out_z_list", [mult_relin(xs[0], xs[1], "z_0"), mult_relin(xs[0], xs[2], "z_1")])])

    set_fhe_param(bfv16())
    ins = [Ct(f"x_{i}", level=lv) for i in range(4)]
    x = ins
    for i in range(2):
        x = [mult_relin(x[j], x[(j + 1) % 4], f"y_{i * 4 + j}") for j in range(4)]
    emit("bfv16k_braid", [Argument("in_list", ins)], [Argument("out_list", x)])

    set_fhe_param(bfv16())
    xs = [Ct(f"x_{i}", level=lv) for i in range(4)]
    co = [Ct(f"coeffs_{i}", level=lv) for i in range(3)]
    zs = []
    for i in range(4):
        ax = mult_relin(xs[i], co[0], f"ax_{i}")
        ax2 = mult_relin(ax, xs[i], f"ax2_{i}")
        bx = mult_relin(co[1], xs[i], f"bx_{i}")
        zs.append(add(add(ax2, bx, f"ax2_bx_{i}"), co[2], f"ax2_bx_c_{i}"))
    emit("bfv16k_poly", [Argument("in_x_list", xs), Argument("in_a_list", co)], [Argument("out_z_list", zs)])

    # CKKS ct x pt multiply-accumulate with NTT-domain plaintexts (test_gpu_ckks.py:233: m = 2..20)
    for m in (2, 5, 20):
        set_fhe_param(CkksParam.create_default_param(n=16384))
        cs = [CkksCiphertextNode(f"c_{i}", level=2) for i in range(m)]
        ps = [CkksPlaintextNode(f"p_{i}", level=2) for i in range(m)]
        z = ct_pt_mult_accumulate(cs, ps)
        emit(f"ckks16k_ct_pt_mac_m{m}", [Argument("in_c_list", cs), Argument("in_p_list", ps)], [Argument("out_z_list", [z])])

    # default-parameter (n = 16384) level sweeps of the two hot-path cases
    for lv in (1, 3, 5):
        set_fhe_param(CkksParam.create_default_param(n=16384))
        xs = [CkksCiphertextNode(f"x_{i}", level=lv) for i in range(2)]
        ys = [CkksCiphertextNode(f"y_{i}", level=lv) for i in range(2)]
        zs = [rescale(mult_relin(xs[i], ys[i]), f"z_{i}") for i in range(2)]
        emit(f"ckks16k_cmc_relin_rescale_l{lv}", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", zs)])
        set_fhe_param(CkksParam.create_default_param(n=16384))
        xs = [CkksCiphertextNode(f"x_{i}", level=lv) for i in range(2)]
        ys = [rotate_cols(xs[i], [1, 3]) for i in range(2)]
        emit(f"ckks16k_rotate_col_l{lv}", [Argument("in_x_list", xs)], [Argument("out_y_list", ys)])
    for lv in (1, 3, 5):
        set_fhe_param(BfvParam.create_default_param(n=16384))
        xs = [BfvCiphertextNode(f"x_{i}", level=lv) for i in range(2)]
        ys = [BfvCiphertextNode(f"y_{i}", level=lv) for i in range(2)]
        zs = [rescale(mult_relin(xs[i], ys[i]), f"z_{i}") for i in range(2)]
        emit(f"bfv16k_cmc_relin_rescale_l{lv}", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", zs)])
        set_fhe_param(BfvParam.create_default_param(n=16384))
        xs = [BfvCiphertextNode(f"x_{i}", level=lv) for i in range(2)]
        ys = [rotate_cols(xs[i], [1, 3]) for i in range(2)]
        emit(f"bfv16k_rotate_col_l{lv}", [Argument("in_x_list", xs)], [Argument("out_y_list", ys)])

    # sparse-slot CKKS parameter set (test_gpu_ckks.py:47-48): slots only matter to encode/decode
    set_fhe_param(ckks8k())
    p = CkksParam.create_custom_param(n=CKKS8K["n"], q=list(CKKS8K["q"]), p=list(CKKS8K["p"]), scale=CKKS8K["scale"])
    p.set_slots(2048)
    set_fhe_param(p)
    xs = [CkksCiphertextNode(f"x_{i}", level=4) for i in range(2)]
    ys = [CkksCiphertextNode(f"y_{i}", level=4) for i in range(2)]
    zs = [rescale(mult_relin(xs[i], ys[i]), f"z_{i}") for i in range(2)]
    emit("ckks8k_sparse_cmc_relin_rescale", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", zs)])

    # custom node at the end of the graph (test_gpu_bfv.py:650-676), mapped onto the harness' "passthrough" executor
    set_fhe_param(bfv8k())
    xs = [BfvCiphertextNode(f"x_{i}", level=2) for i in range(N_OP)]
    ys = [BfvCiphertextNode(f"y_{i}", level=2) for i in range(N_OP)]
    ms = [relin(mult(xs[i], ys[i], f"x_mult_y_{i}"), f"m_{i}") for i in range(N_OP)]
    zs = []
    for i in range(N_OP):
        zi = BfvCiphertextNode(f"z_{i}", level=2)
        custom_compute(inputs=[ms[i]], output=zi, type="passthrough")
        zs.append(zi)
    emit("bfv8k_custom_at_end", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", zs)])


def gen_custom_start():
    """Custom (host) nodes at the START of the graph feeding device nodes: unittests/test_gpu_bfv.py:570-647 (custom_cmpac /
    custom_compute_at_start): custom data inputs -> `encode_ringt` / `encode` custom nodes -> plaintexts the device consumes."""
    for tag, mkparam, Ct, PtR, Pt, lv in (("bfv8k", bfv8k, BfvCiphertextNode, BfvPlaintextRingtNode, BfvPlaintextNode, 2),
                                          ("ckks8k", ckks8k, CkksCiphertextNode, CkksPlaintextRingtNode, CkksPlaintextNode, 4)):
        set_fhe_param(mkparam())
        xs = [Ct(f"x_{i}", level=lv) for i in range(7)]
        ys = [CustomDataNode(type="msg", id=f"y_{i}", attributes={"level": 0}) for i in range(7)]
        ys.append(CustomDataNode("y_7"))
        prods = []
        for i in range(7):
            e = PtR(f"encoded_y_ringt_{i}")
            custom_compute(inputs=[ys[i]], output=e, type="encode_ringt")
            prods.append(mult(xs[i], e, f"mult_{i}"))
        res = prods[0]
        for i in range(1, 7):
            res = add(res, prods[i], f"partial_sum_{i}")
        last = Pt("encoded_y_last", level=res.level)
        custom_compute(inputs=[ys[7]], output=last, type="encode", attributes={"level": res.level})
        final = add(res, last, "z_final")
        emit(f"{tag}_custom_cmpac", [Argument("in_x_list", xs), Argument("in_y_list", ys)], [Argument("out_z_list", final)])


def gen_power_mul_coeff(max_power=300, lanes=(2, 1, 2)):
    """unittests/test_gpu_bfv.py:790-880 (BFV_power_mul_coeff): per lane a ct x pt_ringt multiply-accumulate over `max_power`
    ciphertext powers, split into 4 partial sums above 100 terms, + a plaintext constant, rescale; the ciphertexts are shared
    by every lane.  The reference runs it at n = 16384 with max_power 1137 (and skips it when the power-DAG files of its sister
    test are absent); here the same generator at the n = 8192 fixture parameters with 300 powers (75-term sums)."""
    set_fhe_param(bfv8k())
    lv = 1
    all_c = [[BfvCiphertextNode(f"c{k}_{i + 1}", level=lv) for i in range(max_power)] for k in range(lanes[1])]
    all_p0, all_p, all_z = [], [], []
    for r in range(lanes[0]):
        p0_l, p_l, z_l = [], [], []
        for k in range(lanes[1]):
            p0 = [BfvPlaintextNode(f"p{r}_{k}_{i}_0", level=lv) for i in range(lanes[2])]
            pl = [[BfvPlaintextRingtNode(f"p{r}_{k}_{i}_{j + 1}") for j in range(max_power)] for i in range(lanes[2])]
            p0_l.append(p0)
            p_l.append(pl)
            z = []
            for i in range(lanes[2]):
                n_split, x_mac = 4, None
                for i_split in range(n_split):
                    a, b = int(max_power / n_split * i_split), int(max_power / n_split * (i_split + 1))
                    x = ct_pt_mult_accumulate(all_c[k][a:b], pl[i][a:b])
                    x_mac = x if i_split == 0 else add(x, x_mac)
                y = add(x_mac, p0[i], f"y{r}_{k}_{i}")
                z.append(rescale(y, f"z{r}_{k}_{i}"))
            z_l.append(z)
        all_p0.append(p0_l)
        all_p.append(p_l)
        all_z.append(z_l)
    emit("bfv8k_power_mul_coeff", [Argument("in_c_list", all_c), Argument("in_p0_list", all_p0), Argument("in_p_list", all_p)],
         [Argument("out_z_list", all_z)])


def gen_euclidean_distance(n_ct=2, pack=4, skip=256, level=3):
    """examples/ckks_euclidean_distance_cpu/ckks_euclidean_distance_cpu.py:26-70 at the default n = 16384 set: add, square +
    relin + rescale, rotate_cols by skip, 2*skip, 3*skip (the last one two key switches through the power-of-two keys), sums,
    a pt_ringt mask, rescale.  The example uses one ciphertext pair; two here so that the outer sum has a term."""
    set_fhe_param(CkksParam.create_default_param(n=16384))  # the example names a constructor that no longer exists
    xs = [CkksCiphertextNode(f"x_{i}", level=level) for i in range(n_ct)]
    ws = [CkksCiphertextNode(f"w_{i}", level=level) for i in range(n_ct)]
    mask = CkksPlaintextRingtNode(id="mask")
    sums = []
    for i in range(n_ct):
        z = add(xs[i], ws[i], f"z_{i}")
        u = rescale(mult_relin(z, z), f"u_{i}")
        rots, step = [u], skip
        for j in range(pack - 1):
            rots.append(rotate_cols(u, step, f"u_rot_{i}_{j}")[0])
            step += skip
        acc = rots[0]
        for k in range(1, len(rots)):
            acc = add(acc, rots[k], f"sum_{i}_{k}")
        sums.append(acc)
    total = sums[0]
    for i in range(1, n_ct):
        total = add(total, sums[i], f"sum_{i}")
    distance = rescale(mult(total, mask, "distance"))
    emit("ckks16k_euclidean_distance", [Argument("x_input", xs), Argument("w_input_inv", ws), Argument("mask", mask)],
         [Argument("d", distance)])


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if sys.argv[1:] == ["power"]:
        gen_power_mul_coeff()
        sys.exit(0)
    if sys.argv[1:] == ["euclid"]:
        gen_euclidean_distance()
        sys.exit(0)
    if sys.argv[1:] == ["custom"]:
        gen_custom_start()
        sys.exit(0)
    if sys.argv[1:] == ["btp"]:
        gen_bootstrap()
        sys.exit(0)
    if sys.argv[1:] == ["extras"]:
        gen_reference_extras()
        sys.exit(0)
    if sys.argv[1:] == ["cfg3x"]:
        gen_cfg3_replicas(8)
        gen_cfg3_replicas(64)  # throughput fixture for tools/multi_device_run.py; gzip it after generation
        sys.exit(0)
    if sys.argv[1:] == ["seal"]:
        gen_seal_flavoured()
        sys.exit(0)
    if sys.argv[1:] == ["n65536"]:
        gen_logreg_n65536()
        sys.exit(0)
    if sys.argv[1:] == ["conv"]:
        gen_conv("conv_64x64_5x5", (64, 64), (5, 5))
        gen_conv("conv_4x4_5x5_stress", (4, 4), (5, 5))
        sys.exit(0)
    gen_scheme("bfv8k", bfv8k, BfvCiphertextNode, BfvPlaintextNode, BfvPlaintextRingtNode, BfvPlaintextMulNode, 2, 1)
    gen_scheme("ckks8k", ckks8k, CkksCiphertextNode, CkksPlaintextNode, CkksPlaintextRingtNode, CkksPlaintextMulNode, 4, 1)
    gen_ckks_extra("ckks8k", ckks8k, 4)
    gen_configs()
