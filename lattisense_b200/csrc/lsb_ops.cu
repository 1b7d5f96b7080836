// Ciphertext-level operations: host orchestration of the kernel pipeline (no arithmetic here).
//
// Hybrid key-switch (relinearize / rotate) per batch chunk, l = level+1, k = |P|, beta = ceil(l/k):
//   [INTT c] -> ModUp (beta digits, Lattigo float-v CRT) -> NTT of the extended limbs ->
//   inner product with the evaluation key -> INTT of the 2k P-limbs -> ModDown extension ->
//   NTT -> (acc - ext) * P^-1 (+ c0/c1).
// A batch is processed in chunks bounded by the scratch budget (LSB_SCRATCH_MB, default 8 GiB): measured
// on B200 (profiles/r01_scratch_sweep.txt) large launches beat L2-resident small chunks 2:1 because
// the kernels are integer-pipe bound and small launches leave SMs idle at the tail.
#include <cstdlib>
#include <algorithm>
#include "lsb_ops.h"
#include "lsb_host_math.h"

namespace lsb {

static BatchPtr adv(BatchPtr p, int b0) {
    if (p.table) p.table += b0;
    else p.base += (size_t)b0 * p.stride;
    return p;
}

static size_t scratch_budget_words() {
    static size_t v = [] {
        const char* e = getenv("LSB_SCRATCH_MB");
        size_t mb = e ? (size_t)atol(e) : 8192;
        if (mb < 8) mb = 8;
        return mb * (1u << 20) / 8;
    }();
    return v;
}

// Largest chunk within the scratch budget whose dominant NTT launch fills whole waves of SMs.
static int pick_chunk(const OpEnv& e, size_t item_words, int batch, int rows_per_item) {
    if (e.chunk > 0) return std::min(e.chunk, batch);
    int cmax = (int)std::max<size_t>(1, scratch_budget_words() / std::max<size_t>(item_words, 1));
    if (cmax >= batch) return batch;
    if (e.c.logN > 14 || rows_per_item <= 0) return cmax;
    const int sms = e.c.dt.sm_count;
    int best = cmax;
    double best_eff = 0;
    for (int c = cmax; c >= std::max(1, cmax / 2); c--) {
        const int rows = c * rows_per_item;
        const double eff = (double)rows / (double)(((rows + sms - 1) / sms) * sms);
        if (eff > best_eff + 1e-9) {
            best_eff = eff;
            best = c;
        }
    }
    return best;
}

static RowMap rowmap_seq(int n, int in0, int out0, int mod0, int period) {
    LSB_CHECK(n <= kMaxRows, "lsb: too many rows in one launch");
    RowMap m;
    m.n = n;
    for (int r = 0; r < n; r++) {
        m.in_slot[r] = (uint8_t)(in0 + r);
        m.out_slot[r] = (uint8_t)(out0 + r);
        m.mod[r] = (uint8_t)(mod0 + (period ? r % period : r));
    }
    return m;
}

// ---------------------------------------------------------------------------------------------
// key-switch core
// ---------------------------------------------------------------------------------------------
static int ks_scratch_limbs(const Context& c, int level) {
    const int l = level + 1, lk = l + c.np;
    return c.beta(level) * lk + 2 * lk + 2 * l;
}

// The key-switch in two halves, so that several rotations of ONE ciphertext share the first ("hoisting": the CKKS rotation
// key-switches the un-permuted c1 and permutes afterwards, so the decomposition does not depend on the Galois element):
//   ks_decompose: ModUp of the beta digits + NTT of the extended limbs            -> ext [beta][l+k]
//   ks_apply:     inner product with the key, INTT of the P limbs, ModDown, NTT + finish
// res[p*res_ps + i] = (acc_p[i] - md_p[i]) * P^-1 (+ addend[p*l + i] if addmask bit p), NTT domain
static void ks_decompose(OpEnv& e, int level, BatchPtr c_coeff, int coeff_slot0, BatchPtr ext, int nb) {
    Context& c = e.c;
    const DeviceTables& dt = c.dt;
    const int l = level + 1, k = c.np, lk = l + k, beta = c.beta(level), alpha = k;
    // digits go in launch groups whose rows fit the uint8_t slots of one launch (one group for every BASELINE configuration)
    const int grp = c.ks_digit_group(level);
    LSB_CHECK(grp == beta || ext.table == nullptr, "lsb: grouped key-switch digits need a strided scratch block");
    const BexJob* jobs = c.modup_jobs(level, coeff_slot0);
    const BexMeta* meta = c.bex_meta(jobs);
    for (int d0 = 0; d0 < beta; d0 += grp) {
        const int nd = std::min(grp, beta - d0);
        BatchPtr ext_g = ext;
        if (d0) ext_g.base += (size_t)d0 * lk * c.N;
        launch_basis_extend(dt, jobs + d0, nd, c_coeff, ext_g, ext_g, nb, e.s, meta ? meta + d0 : nullptr);
        RowMap m;
        m.n = 0;
        for (int d = d0; d < d0 + nd; d++)
            for (int t = 0; t < lk; t++) {
                if (t < l && t >= d * alpha && t < (d + 1) * alpha) continue;
                LSB_CHECK(m.n < kMaxRows, "lsb: too many extended limbs");
                m.in_slot[m.n] = m.out_slot[m.n] = (uint8_t)((d - d0) * lk + t);
                m.mod[m.n] = (uint8_t)(t < l ? t : c.idxP(t - l));
                m.n++;
            }
        launch_ntt(dt, ext_g, ext_g, m, nb, false, e.s);
    }
}
// scratch: [acc 2(l+k) | md 2l] limbs per item
static void ks_apply(OpEnv& e, int level, BatchPtr ext, BatchPtr c_ntt, int ntt_slot0, const u64* evk, int key_level, BatchPtr addend,
                     int addmask, BatchPtr res, int res_ps, u64* scratch, size_t item_words, int nb) {
    Context& c = e.c;
    const DeviceTables& dt = c.dt;
    const int l = level + 1, k = c.np, lk = l + k, beta = c.beta(level), alpha = k;
    const size_t N = c.N;
    LSB_CHECK(key_level >= level, "lsb: evaluation key level below ciphertext level");
    BatchPtr acc = bp_strided(scratch, item_words);
    BatchPtr md = bp_strided(scratch + (size_t)(2 * lk) * N, item_words);
    KsInnerArgs ia;
    ia.ext = ext;
    ia.cntt = c_ntt;
    ia.cntt_slot0 = ntt_slot0;
    ia.evk = evk;
    ia.acc = acc;
    ia.l = l;
    ia.k = k;
    ia.beta = beta;
    ia.alpha = alpha;
    ia.key_level = key_level;
    ia.nq = c.nq;
    ia.mapped = 0;
    ia.rows = 0;
    {
        u64 qmax = 0;
        for (int i = 0; i < l; i++) qmax = std::max(qmax, c.moduli[i]);
        for (int j = 0; j < k; j++) qmax = std::max(qmax, c.moduli[c.idxP(j)]);
        ia.fold = (int)std::min<u64>(1u << 20, ((1ull << 63) - 1) / qmax);
        static const bool elem_f64 = [] {
            const char* e = getenv("LSB_ELEM_F64");
            return !(e && atoi(e) == 0);
        }();
        ia.f64 = (elem_f64 && qmax < kElemF64Limit) ? 1 : 0;
    }
    launch_ks_inner(dt, ia, nb, e.s);
    {
        RowMap m;
        m.n = 0;
        for (int p = 0; p < 2; p++)
            for (int j = 0; j < k; j++) {
                m.in_slot[m.n] = m.out_slot[m.n] = (uint8_t)(p * lk + l + j);
                m.mod[m.n] = (uint8_t)c.idxP(j);
                m.n++;
            }
        launch_ntt(dt, acc, acc, m, nb, true, e.s);
    }
    {
        const BexJob* jobs = c.moddown_jobs(level);
        launch_basis_extend(dt, jobs, 2, acc, md, md, nb, e.s, c.bex_meta(jobs));
    }
    {
        const FinishMap* fm = c.moddown_finish(level, addmask, res_ps);
        NttFinish fin{fm, c.finish_host(fm), acc, addend, res, addmask != 0};
        launch_ntt_finish(dt, md, fin, nb, e.s);
    }
}
static void ks_core(OpEnv& e, int level, BatchPtr c_coeff, int coeff_slot0, BatchPtr c_ntt, int ntt_slot0,
                    const u64* evk, int key_level, BatchPtr addend, int addmask, BatchPtr res, int res_ps,
                    u64* scratch, size_t item_words, int nb) {
    const int l = level + 1, lk = l + e.c.np, beta = e.c.beta(level);
    BatchPtr ext = bp_strided(scratch, item_words);
    ks_decompose(e, level, c_coeff, coeff_slot0, ext, nb);
    ks_apply(e, level, ext, c_ntt, ntt_slot0, evk, key_level, addend, addmask, res, res_ps, scratch + (size_t)beta * lk * e.c.N, item_words, nb);
}

// ---------------------------------------------------------------------------------------------
// mult
// ---------------------------------------------------------------------------------------------
static void ckks_tensor(OpEnv& e, int level, BatchPtr a, BatchPtr b, BatchPtr out3, int nb) {
    const int l = level + 1;
    uint8_t mods[64];
    LSB_CHECK(l <= 64, "lsb: too many limbs");
    for (int i = 0; i < l; i++) mods[i] = (uint8_t)i;
    launch_tensor(e.c.dt, a, b, out3, l, l, l, mods, nb, e.s);
}

static int bfv_mult_scratch_limbs(const Context& c, int level) {
    const int l = level + 1, lm = c.level_qmul[level] + 1, L = l + lm;
    return 2 * L + 2 * L + 3 * L + 3 * lm;
}
// bfv evaluator tensorAndRescale (oracle/orc_core.hpp:420-484)
static void bfv_mult_impl(OpEnv& e, int level, BatchPtr a, BatchPtr b, BatchPtr out3, u64* scratch, size_t item_words,
                          int nb) {
    Context& c = e.c;
    const DeviceTables& dt = c.dt;
    const int l = level + 1, lm = c.level_qmul[level] + 1, L = l + lm;
    const size_t N = c.N;
    BatchPtr ea = bp_strided(scratch, item_words);
    BatchPtr eb = bp_strided(scratch + (size_t)2 * L * N, item_words);
    BatchPtr prod = bp_strided(scratch + (size_t)4 * L * N, item_words);
    BatchPtr buf = bp_strided(scratch + (size_t)7 * L * N, item_words);
    const BexJob* up = c.bfv_q_to_qmul_jobs(level);
    RowMap mq, mm;  // Q limbs: input -> extended block ; QMul limbs: in place
    mq.n = mm.n = 0;
    for (int p = 0; p < 2; p++) {
        for (int i = 0; i < l; i++) {
            mq.in_slot[mq.n] = (uint8_t)(p * l + i);
            mq.out_slot[mq.n] = (uint8_t)(p * L + i);
            mq.mod[mq.n++] = (uint8_t)i;
        }
        for (int j = 0; j < lm; j++) {
            mm.in_slot[mm.n] = mm.out_slot[mm.n] = (uint8_t)(p * L + l + j);
            mm.mod[mm.n++] = (uint8_t)c.idxQM(j);
        }
    }
    const BatchPtr ins[2] = {a, b};
    const BatchPtr exts[2] = {ea, eb};
    for (int w = 0; w < 2; w++) {
        launch_basis_extend(dt, up, 2, ins[w], exts[w], exts[w], nb, e.s);
        launch_ntt(dt, ins[w], exts[w], mq, nb, false, e.s);
        launch_ntt(dt, exts[w], exts[w], mm, nb, false, e.s);
    }
    uint8_t mods[64];
    LSB_CHECK(L <= 64, "lsb: too many limbs");
    for (int i = 0; i < l; i++) mods[i] = (uint8_t)i;
    for (int j = 0; j < lm; j++) mods[l + j] = (uint8_t)c.idxQM(j);
    launch_tensor(dt, ea, eb, prod, L, L, L, mods, nb, e.s);
    {
        RowMap m;
        m.n = 3 * L;
        LSB_CHECK(m.n <= kMaxRows, "lsb: too many limbs");
        for (int r = 0; r < 3 * L; r++) {
            m.in_slot[r] = m.out_slot[r] = (uint8_t)r;
            m.mod[r] = mods[r % L];
        }
        launch_ntt(dt, prod, prod, m, nb, true, e.s);
    }
    launch_basis_extend(dt, c.bfv_quant1_jobs(level), 3, prod, buf, prod, nb, e.s);
    launch_basis_extend(dt, c.bfv_quant2_jobs(level), 3, buf, out3, buf, nb, e.s);
}

void op_mult(OpEnv& e, int level, BatchPtr a, BatchPtr b, BatchPtr out3, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    if (e.c.scheme == SCHEME_CKKS) {
        ckks_tensor(e, level, a, b, out3, batch);
        return;
    }
    const size_t item = (size_t)bfv_mult_scratch_limbs(e.c, level) * e.c.N;
    const int l = level + 1, L = l + e.c.level_qmul[level] + 1;
    const int chunk = pick_chunk(e, item, batch, 3 * L);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        bfv_mult_impl(e, level, adv(a, b0), adv(b, b0), adv(out3, b0), scratch, item, nb);
    }
}

// ---------------------------------------------------------------------------------------------
// relinearize
// ---------------------------------------------------------------------------------------------
static int relin_scratch_limbs(const Context& c, int level) {
    const int l = level + 1;
    return (c.scheme == SCHEME_CKKS ? l : 3 * l) + ks_scratch_limbs(c, level);
}
static void relin_impl(OpEnv& e, int level, BatchPtr ct3, const u64* rlk, int key_level, BatchPtr out2, u64* scratch,
                       size_t item_words, int nb) {
    Context& c = e.c;
    const DeviceTables& dt = c.dt;
    const int l = level + 1;
    const size_t N = c.N;
    if (c.scheme == SCHEME_CKKS) {
        BatchPtr c2 = bp_strided(scratch, item_words);
        launch_ntt(dt, ct3, c2, rowmap_seq(l, 2 * l, 0, 0, l), nb, true, e.s);
        ks_core(e, level, c2, 0, ct3, 2 * l, rlk, key_level, ct3, 3, out2, l, scratch + (size_t)l * N, item_words, nb);
    } else {
        // coefficient-domain ciphertext: c2 itself feeds ModUp, NTT(c2) feeds the own-digit products
        BatchPtr cn = bp_strided(scratch, item_words);
        BatchPtr tmp = bp_strided(scratch + (size_t)l * N, item_words);
        launch_ntt(dt, ct3, cn, rowmap_seq(l, 2 * l, 0, 0, l), nb, false, e.s);
        ks_core(e, level, ct3, 2 * l, cn, 0, rlk, key_level, tmp, 0, tmp, l, scratch + (size_t)3 * l * N, item_words,
                nb);
        launch_ntt(dt, tmp, tmp, rowmap_seq(2 * l, 0, 0, 0, l), nb, true, e.s);
        launch_binary(dt, OP_ADD, tmp, ct3, out2, rowmap_seq(2 * l, 0, 0, 0, l), 2 * l, nb, e.s);
    }
}
void op_relin(OpEnv& e, int level, BatchPtr ct3, const u64* rlk, int key_level, BatchPtr out2, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    const size_t item = (size_t)relin_scratch_limbs(e.c, level) * e.c.N;
    const int l = level + 1;
    const int chunk = pick_chunk(e, item, batch, e.c.beta(level) * (l + e.c.np) - l);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        relin_impl(e, level, adv(ct3, b0), rlk, key_level, adv(out2, b0), scratch, item, nb);
    }
}

// ---------------------------------------------------------------------------------------------
// rescale / drop level
// ---------------------------------------------------------------------------------------------
static int rescale_scratch_limbs(const Context& c, int level, int npoly) {
    return c.scheme == SCHEME_CKKS ? npoly * (1 + level) : 0;
}
static void rescale_impl(OpEnv& e, int level, int npoly, BatchPtr in, BatchPtr out, u64* scratch, size_t item_words,
                         int nb) {
    Context& c = e.c;
    const DeviceTables& dt = c.dt;
    const int l = level + 1;
    const size_t N = c.N;
    LSB_CHECK(level >= 1, "lsb: cannot rescale a level-0 ciphertext");
    if (c.scheme == SCHEME_BFV) {
        launch_bfv_rescale(dt, in, out, npoly, level, c.qlinv(level), nb, e.s);
        return;
    }
    // ring.DivRoundByLastModulusNTT
    BatchPtr last = bp_strided(scratch, item_words);
    BatchPtr tmp = bp_strided(scratch + (size_t)npoly * N, item_words);
    RowMap m;
    m.n = npoly;
    for (int p = 0; p < npoly; p++) {
        m.in_slot[p] = (uint8_t)(p * l + level);
        m.out_slot[p] = (uint8_t)p;
        m.mod[p] = (uint8_t)level;
    }
    launch_ntt(dt, in, last, m, nb, true, e.s);
    launch_rescale_lift(dt, last, 1, tmp, npoly, level, nb, e.s);
    {
        const FinishMap* fm = c.rescale_finish(level, npoly);
        NttFinish fin{fm, c.finish_host(fm), in, tmp, out, false};
        launch_ntt_finish(dt, tmp, fin, nb, e.s);
    }
}
void op_rescale(OpEnv& e, int level, int npoly, BatchPtr in, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    const size_t item = (size_t)rescale_scratch_limbs(e.c, level, npoly) * e.c.N;
    const int chunk = item ? pick_chunk(e, item, batch, npoly * level) : batch;
    u64* scratch = item ? e.ws.reserve(item * chunk, e.s) : nullptr;
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        rescale_impl(e, level, npoly, adv(in, b0), adv(out, b0), scratch, item, nb);
    }
}

// CKKS ModRaise (first step of bootstrapping; Lattigo ckks/bootstrapping `modUpFromQ0`): a level-0 ciphertext
// [npoly][1][N] (NTT domain) is re-read modulo q_0..q_top -- INTT of the q_0 limb, centred lift to every other prime,
// NTT of all limbs -- and comes out at level `top` encrypting Delta*m + q_0*I.
void op_mod_raise(OpEnv& e, int top, int npoly, BatchPtr in, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    Context& c = e.c;
    LSB_CHECK(c.scheme == SCHEME_CKKS, "lsb: mod_raise is a CKKS operation");
    LSB_CHECK(top >= 1 && top < c.nq && npoly >= 1 && npoly * (top + 1) <= kMaxRows, "lsb: mod_raise: bad level");
    const int l = top + 1;
    const size_t item = (size_t)npoly * c.N;
    const int chunk = pick_chunk(e, item, batch, npoly * l);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        BatchPtr last = bp_strided(scratch, item);
        launch_ntt(c.dt, adv(in, b0), last, rowmap_seq(npoly, 0, 0, 0, 1), nb, true, e.s);
        launch_rescale_lift(c.dt, last, 1, adv(out, b0), npoly, top, nb, e.s, true);
        launch_ntt(c.dt, adv(out, b0), adv(out, b0), rowmap_seq(npoly * l, 0, 0, 0, l), nb, false, e.s);
    }
}

// Diagonal-form linear transform, baby-step / giant-step (the CoeffsToSlots / SlotsToCoeffs matrices of bootstrapping;
// frontend/bootstrap_params.py:159-209 fixes the split and the rotation set):
//   out = rescale( sum_g rot( sum_b pt_{g,b} * rot(ct, b), g ) ),   diagonal k = g + b, b = k mod n1,
// with the plaintexts already rotated by -g and in the NTT domain at `level`.  Every inner sum is one multiply-accumulate
// pass (k_sum_terms), the outer sum another; rotations are the ordinary key switches (hoisting is the next step).
void op_linear_transform(OpEnv& e, int level, const u64* ct, int n_diag, const int* diag_rot, const u64* const* diag_pt, int n1,
                         int n_keys, const int* key_rot, const u64* const* glk, int key_level, u64* out) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    Context& c = e.c;
    LSB_CHECK(c.scheme == SCHEME_CKKS && level >= 1 && level < c.nq && n_diag >= 1 && n1 >= 1, "lsb: linear_transform: bad arguments");
    const int slots = c.N / 2, l = level + 1;
    const size_t words = (size_t)2 * l * c.N;
    auto key_for = [&](int rot) -> const u64* {
        for (int i = 0; i < n_keys; i++)
            if (key_rot[i] == rot) return glk[i];
        throw std::runtime_error("lsb: linear_transform: no Galois key for rotation " + std::to_string(rot));
    };
    auto gal_of = [&](int rot) {
        const u64 m = 2ull * c.N;
        u64 r = 1, b = 5 % m;
        for (int x = rot; x; x >>= 1, b = b * b % m)
            if (x & 1) r = r * b % m;
        return r;
    };
    std::vector<int> babies, giants;
    for (int i = 0; i < n_diag; i++) {
        const int k = ((diag_rot[i] % slots) + slots) % slots, b = k % n1, g = k - b;
        if (std::find(babies.begin(), babies.end(), b) == babies.end()) babies.push_back(b);
        if (std::find(giants.begin(), giants.end(), g) == giants.end()) giants.push_back(g);
    }
    std::sort(giants.begin(), giants.end());
    // device buffers: rotated inputs, one accumulator per giant step, pointer tables
    u64* pool = nullptr;
    const size_t n_buf = babies.size() + 2 * giants.size() + 1;
    LSB_CUDA(cudaMallocAsync((void**)&pool, n_buf * words * sizeof(u64), e.s));
    const u64** tab = nullptr;
    const size_t tab_len = 2 * (size_t)n_diag + 2 * giants.size() + 2;
    LSB_CUDA(cudaMallocAsync((void**)&tab, tab_len * sizeof(u64*), e.s));
    size_t next = 0;
    auto take = [&] { return pool + (next++) * words; };
    std::vector<const u64*> baby_ptr(babies.size());
    for (size_t i = 0; i < babies.size(); i++) {
        if (babies[i] == 0) {
            baby_ptr[i] = ct;
            continue;
        }
        u64* dst = take();
        op_rotate(e, level, bp_strided(ct, words), key_for(babies[i]), key_level, gal_of(babies[i]), bp_strided(dst, words), 1);
        baby_ptr[i] = dst;
    }
    std::vector<const u64*> host_tab;
    std::vector<const u64*> giant_out;
    size_t tab_used = 0;
    auto upload = [&](const std::vector<const u64*>& v) {
        const u64** d = tab + tab_used;
        LSB_CUDA(cudaMemcpyAsync(d, v.data(), v.size() * sizeof(u64*), cudaMemcpyHostToDevice, e.s));
        tab_used += v.size();
        return d;
    };
    for (int g : giants) {
        std::vector<const u64*> cts, pts;
        for (int i = 0; i < n_diag; i++) {
            const int k = ((diag_rot[i] % slots) + slots) % slots;
            if (k - k % n1 != g) continue;
            const size_t bi = std::find(babies.begin(), babies.end(), k % n1) - babies.begin();
            cts.push_back(baby_ptr[bi]);
            pts.push_back(diag_pt[i]);
        }
        u64* acc = take();
        std::vector<const u64*> o{acc};
        const u64** d_ct = upload(cts);
        const u64** d_pt = upload(pts);
        const u64** d_o = upload(o);
        launch_sum_terms(c.dt, d_ct, d_pt, (u64* const*)d_o, 2, l, (int)cts.size(), 0, 1, e.s);
        if (g == 0) {
            giant_out.push_back(acc);
        } else {
            u64* r = take();
            op_rotate(e, level, bp_strided(acc, words), key_for(g), key_level, gal_of(g), bp_strided(r, words), 1);
            giant_out.push_back(r);
        }
    }
    u64* total = take();
    {
        std::vector<const u64*> o{total};
        const u64** d_ct = upload(giant_out);
        const u64** d_o = upload(o);
        launch_sum_terms(c.dt, d_ct, nullptr, (u64* const*)d_o, 2, l, 0, (int)giant_out.size(), 1, e.s);
    }
    LSB_CHECK(tab_used <= tab_len && next <= n_buf, "lsb: linear_transform: internal buffer accounting");
    op_rescale(e, level, 2, bp_strided(total, words), bp_strided(out, (size_t)2 * level * c.N), 1);
    LSB_CUDA(cudaFreeAsync(pool, e.s));
    LSB_CUDA(cudaFreeAsync(tab, e.s));
}

void op_drop_level(OpEnv& e, int level, int npoly, BatchPtr in, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    LSB_CHECK(level >= 1, "lsb: cannot drop below level 0");
    const int l = level + 1;
    RowMap m;
    m.n = npoly * (l - 1);
    LSB_CHECK(m.n <= kMaxRows, "lsb: too many limbs");
    for (int p = 0; p < npoly; p++)
        for (int i = 0; i < l - 1; i++) {
            const int r = p * (l - 1) + i;
            m.in_slot[r] = (uint8_t)(p * l + i);
            m.out_slot[r] = (uint8_t)r;
            m.mod[r] = (uint8_t)i;
        }
    // copy through the binary kernel with no second operand rows
    launch_binary(e.c.dt, OP_ADD, in, in, out, m, 0, batch, e.s);
}

// ---------------------------------------------------------------------------------------------
// rotate (automorphism + key-switch)
// ---------------------------------------------------------------------------------------------
static int rotate_scratch_limbs(const Context& c, int level) { return 3 * (level + 1) + ks_scratch_limbs(c, level); }
static int dual_stream_min_batch();
static void rotate_impl(OpEnv& e, int level, BatchPtr ct, const u64* glk, int key_level, u64 gal, BatchPtr out,
                        u64* scratch, size_t item_words, int nb, cudaEvent_t after_first = nullptr) {
    Context& c = e.c;
    const DeviceTables& dt = c.dt;
    const int l = level + 1;
    const size_t N = c.N;
    BatchPtr c1x = bp_strided(scratch, item_words);                   // c1 in the other domain
    BatchPtr tmp = bp_strided(scratch + (size_t)l * N, item_words);   // [2][l]
    u64* ks_scratch = scratch + (size_t)3 * l * N;
    if (c.scheme == SCHEME_CKKS) {
        // ckks evaluator permuteNTT: key-switch c1, add c0, permute both (oracle/orc_core.hpp:402-414)
        launch_ntt(dt, ct, c1x, rowmap_seq(l, l, 0, 0, l), nb, true, e.s);
        if (after_first) LSB_CUDA(cudaEventRecord(after_first, e.s));
        ks_core(e, level, c1x, 0, ct, l, glk, key_level, ct, 1, tmp, l, ks_scratch, item_words, nb);
        launch_permute_ntt(dt, tmp, out, rowmap_seq(2 * l, 0, 0, 0, l), c.permute_index(gal), nb, e.s);
    } else {
        launch_ntt(dt, ct, c1x, rowmap_seq(l, l, 0, 0, l), nb, false, e.s);
        ks_core(e, level, ct, l, c1x, 0, glk, key_level, tmp, 0, tmp, l, ks_scratch, item_words, nb);
        launch_ntt(dt, tmp, tmp, rowmap_seq(2 * l, 0, 0, 0, l), nb, true, e.s);
        launch_binary(dt, OP_ADD, tmp, ct, tmp, rowmap_seq(2 * l, 0, 0, 0, l), l, nb, e.s);
        launch_permute_coeff(dt, tmp, out, rowmap_seq(2 * l, 0, 0, 0, l), gal, nb, e.s);
    }
}
void op_rotate(OpEnv& e, int level, BatchPtr ct, const u64* glk, int key_level, u64 gal, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    const size_t item = (size_t)rotate_scratch_limbs(e.c, level) * e.c.N;
    const int l = level + 1;
    Context& c = e.c;
    const bool profiling = c.dt.prof && c.dt.prof->enabled.load(std::memory_order_relaxed);
    if (c.scheme == SCHEME_CKKS && e.chunk <= 0 && batch >= dual_stream_min_batch() && !profiling &&
        item * (size_t)batch <= scratch_budget_words()) {
        // large CKKS batches: two half-batches on two streams, one kernel apart, as in mult_relin_chain -- the FP64-bound
        // transforms of one half run beside the HBM-bound inner product / permutation of the other
        const int h0 = (batch + 1) / 2, h1 = batch - h0;
        u64* scratch = e.ws.reserve(item * batch, e.s);
        Context::SidePipe& sp = c.side_pipe();
        OpEnv side{e.c, e.ws, sp.stream, e.chunk};
        rotate_impl(e, level, ct, glk, key_level, gal, out, scratch, item, h0, sp.fork);
        LSB_CUDA(cudaStreamWaitEvent(sp.stream, sp.fork, 0));
        rotate_impl(side, level, adv(ct, h0), glk, key_level, gal, adv(out, h0), scratch + item * h0, item, h1);
        LSB_CUDA(cudaEventRecord(sp.join, sp.stream));
        LSB_CUDA(cudaStreamWaitEvent(e.s, sp.join, 0));
        return;
    }
    const int chunk = pick_chunk(e, item, batch, e.c.beta(level) * (l + e.c.np) - l);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        rotate_impl(e, level, adv(ct, b0), glk, key_level, gal, adv(out, b0), scratch, item, nb);
    }
}

// Key switch without automorphism (CKKS, NTT domain): out = (c0 + KS_0(c1), KS_1(c1)) under `swk` -- the dense <-> sparse
// secret encapsulation around ModRaise in bootstrapping (swk_dts / swk_std, frontend/custom_task.py:1989-1996)
void op_switch_key(OpEnv& e, int level, BatchPtr ct, const u64* swk, int key_level, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    Context& c = e.c;
    LSB_CHECK(c.scheme == SCHEME_CKKS, "lsb: switch_key is a CKKS (NTT-domain) operation");
    const int l = level + 1;
    const size_t N = c.N;
    const size_t item = (size_t)(l + ks_scratch_limbs(c, level)) * N;
    const int chunk = pick_chunk(e, item, batch, c.beta(level) * (l + c.np) - l);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        BatchPtr c1x = bp_strided(scratch, item);
        launch_ntt(c.dt, adv(ct, b0), c1x, rowmap_seq(l, l, 0, 0, l), nb, true, e.s);
        ks_core(e, level, c1x, 0, adv(ct, b0), l, swk, key_level, adv(ct, b0), 1, adv(out, b0), l, scratch + (size_t)l * N, item, nb);
    }
}

// first level_to+1 limbs of every polynomial (drop_level applied level_from - level_to times)
void op_drop_to(OpEnv& e, int level_from, int level_to, int npoly, BatchPtr in, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    LSB_CHECK(level_to >= 0 && level_to <= level_from, "lsb: drop_to: bad levels");
    const int lf = level_from + 1, lt = level_to + 1;
    RowMap m;
    m.n = npoly * lt;
    LSB_CHECK(m.n <= kMaxRows, "lsb: too many limbs");
    for (int p = 0; p < npoly; p++)
        for (int i = 0; i < lt; i++) {
            const int r = p * lt + i;
            m.in_slot[r] = (uint8_t)(p * lf + i);
            m.out_slot[r] = (uint8_t)r;
            m.mod[r] = (uint8_t)i;
        }
    launch_binary(e.c.dt, OP_ADD, in, in, out, m, 0, batch, e.s);
}

// ct times / plus an integer constant given as residues modulo q_0..q_level
void op_scalar(OpEnv& e, int level, int npoly, BatchPtr ct, const u64* residues, bool add, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    const int l = level + 1;
    LSB_CHECK(l <= 64, "lsb: too many limbs");
    Tw consts[64];
    for (int i = 0; i < l; i++) consts[i] = host::make_tw(residues[i] % e.c.moduli[i], e.c.moduli[i]);
    launch_scalar(e.c.dt, ct, out, npoly, l, consts, add, batch, e.s);
}

// n_rot rotations of the SAME ciphertexts with one shared decomposition (CKKS): out_r = rotate(ct, gal_r).  Every limb
// equals op_rotate's (the same kernels on the same data; only the ModUp + forward NTT of the digits is done once).
// Reference: the hoisted baby-step rotations the frontend sketches for BSGS (custom_task.py:1427-1457).
void op_rotate_hoisted(OpEnv& e, int level, BatchPtr ct, int n_rot, const u64* const* glk, int key_level, const u64* gal, const BatchPtr* out,
                       int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    Context& c = e.c;
    LSB_CHECK(c.scheme == SCHEME_CKKS, "lsb: hoisted rotations are a CKKS (NTT-domain) operation");
    const int l = level + 1, lk = l + c.np, beta = c.beta(level);
    const size_t N = c.N;
    // [c1x l | ext beta*lk | tmp 2l | acc 2lk | md 2l]
    const size_t item = (size_t)(l + beta * lk + 2 * l + 2 * lk + 2 * l) * N;
    const int chunk = pick_chunk(e, item, batch, beta * lk - l);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        BatchPtr c1x = bp_strided(scratch, item);
        BatchPtr ext = bp_strided(scratch + (size_t)l * N, item);
        BatchPtr tmp = bp_strided(scratch + (size_t)(l + beta * lk) * N, item);
        u64* rest = scratch + (size_t)(l + beta * lk + 2 * l) * N;
        launch_ntt(c.dt, adv(ct, b0), c1x, rowmap_seq(l, l, 0, 0, l), nb, true, e.s);
        ks_decompose(e, level, c1x, 0, ext, nb);
        for (int r = 0; r < n_rot; r++) {
            ks_apply(e, level, ext, adv(ct, b0), l, glk[r], key_level, adv(ct, b0), 1, tmp, l, rest, item, nb);
            launch_permute_ntt(c.dt, tmp, adv(out[r], b0), rowmap_seq(2 * l, 0, 0, 0, l), c.permute_index(gal[r]), nb, e.s);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// add / sub / neg and plaintext variants
// ---------------------------------------------------------------------------------------------
void op_add(OpEnv& e, int level, int npoly, BatchPtr a, BatchPtr b, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    const int l = level + 1;
    launch_binary(e.c.dt, OP_ADD, a, b, out, rowmap_seq(npoly * l, 0, 0, 0, l), npoly * l, batch, e.s);
}
void op_sub(OpEnv& e, int level, int npoly, BatchPtr a, BatchPtr b, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    const int l = level + 1;
    launch_binary(e.c.dt, OP_SUB, a, b, out, rowmap_seq(npoly * l, 0, 0, 0, l), npoly * l, batch, e.s);
}
void op_neg(OpEnv& e, int level, int npoly, BatchPtr a, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    const int l = level + 1;
    launch_negate(e.c.dt, a, out, rowmap_seq(npoly * l, 0, 0, 0, l), batch, e.s);
}
void op_add_plain(OpEnv& e, int level, int npoly, BatchPtr ct, BatchPtr pt, BatchPtr out, bool subtract, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    const int l = level + 1;
    launch_binary(e.c.dt, subtract ? OP_SUB : OP_ADD, ct, pt, out, rowmap_seq(npoly * l, 0, 0, 0, l), l, batch, e.s);
}
void op_add_plain_ringt(OpEnv& e, int level, int npoly, BatchPtr ct, BatchPtr pt, BatchPtr out, bool subtract,
                        int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    Context& c = e.c;
    const int l = level + 1;
    const size_t item = (size_t)l * c.N;
    const int chunk = pick_chunk(e, item, batch, l);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        BatchPtr lifted = bp_strided(scratch, item);
        if (c.scheme == SCHEME_BFV) {
            launch_bfv_scale_up(c.dt, adv(pt, b0), lifted, l, c.bfv_scale_up_consts(level), nb, e.s);
        } else {
            launch_lift_ringt(c.dt, adv(pt, b0), lifted, l, nb, e.s);
            launch_ntt(c.dt, lifted, lifted, rowmap_seq(l, 0, 0, 0, l), nb, false, e.s);
        }
        launch_binary(c.dt, subtract ? OP_SUB : OP_ADD, adv(ct, b0), lifted, adv(out, b0),
                      rowmap_seq(npoly * l, 0, 0, 0, l), l, nb, e.s);
    }
}
void op_mul_plain(OpEnv& e, int level, int npoly, BatchPtr ct, BatchPtr pt, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    Context& c = e.c;
    const int l = level + 1;
    if (c.scheme == SCHEME_CKKS) {
        launch_mul_plain(c.dt, ct, pt, out, npoly, l, false, batch, e.s);
        return;
    }
    // BFV: ciphertext is in the coefficient domain, pt_mul in the NTT domain
    const size_t item = (size_t)npoly * l * c.N;
    const int chunk = pick_chunk(e, item, batch, npoly * l);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        BatchPtr t = bp_strided(scratch, item);
        launch_ntt(c.dt, adv(ct, b0), t, rowmap_seq(npoly * l, 0, 0, 0, l), nb, false, e.s);
        launch_mul_plain(c.dt, t, adv(pt, b0), t, npoly, l, false, nb, e.s);
        launch_ntt(c.dt, t, adv(out, b0), rowmap_seq(npoly * l, 0, 0, 0, l), nb, true, e.s);
    }
}
void op_mul_plain_ringt(OpEnv& e, int level, int npoly, BatchPtr ct, BatchPtr pt, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    Context& c = e.c;
    const int l = level + 1;
    const size_t item = (size_t)(1 + npoly) * l * c.N;
    const int chunk = pick_chunk(e, item, batch, (1 + npoly) * l);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        BatchPtr ptm = bp_strided(scratch, item);
        BatchPtr t = bp_strided(scratch + (size_t)l * c.N, item);
        launch_lift_ringt(c.dt, adv(pt, b0), ptm, l, nb, e.s);
        launch_ntt(c.dt, ptm, ptm, rowmap_seq(l, 0, 0, 0, l), nb, false, e.s);
        if (c.scheme == SCHEME_CKKS) {
            launch_mul_plain(c.dt, adv(ct, b0), ptm, adv(out, b0), npoly, l, false, nb, e.s);
        } else {
            launch_ntt(c.dt, adv(ct, b0), t, rowmap_seq(npoly * l, 0, 0, 0, l), nb, false, e.s);
            launch_mul_plain(c.dt, t, ptm, t, npoly, l, false, nb, e.s);
            launch_ntt(c.dt, t, adv(out, b0), rowmap_seq(npoly * l, 0, 0, 0, l), nb, true, e.s);
        }
    }
}

// sum_i ct_i x pt_i (+ partial sum): bind_gpu_cmp_sum / bind_gpu_cmpac_sum (mega_ag_executors_gpu.cu:294-408).
// Products are accumulated in the NTT domain; a BFV result takes a single INTT at the end.
void op_mac(OpEnv& e, int level, int nterms, const BatchPtr* cts, const BatchPtr* pts, bool ringt, bool has_acc,
            BatchPtr acc_in, BatchPtr out, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    Context& c = e.c;
    const int l = level + 1;
    const bool bfv = c.scheme == SCHEME_BFV;
    LSB_CHECK(nterms >= 1, "lsb: mac needs at least one term");
    // [acc 2l | ptm l | t 2l]
    const size_t item = (size_t)(2 * l + l + (bfv ? 2 * l : 0)) * c.N;
    const int chunk = pick_chunk(e, item, batch, 2 * l);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) {
        const int nb = std::min(chunk, batch - b0);
        BatchPtr acc = bp_strided(scratch, item);
        BatchPtr ptm = bp_strided(scratch + (size_t)2 * l * c.N, item);
        BatchPtr t = bp_strided(scratch + (size_t)3 * l * c.N, item);
        // a CKKS sum without partial input accumulates straight into the output
        BatchPtr dst = (!bfv && !has_acc) ? adv(out, b0) : acc;
        for (int i = 0; i < nterms; i++) {
            BatchPtr pt = adv(pts[i], b0);
            if (ringt) {
                launch_lift_ringt(c.dt, pt, ptm, l, nb, e.s);
                launch_ntt(c.dt, ptm, ptm, rowmap_seq(l, 0, 0, 0, l), nb, false, e.s);
                pt = ptm;
            }
            BatchPtr ct = adv(cts[i], b0);
            if (bfv) {
                launch_ntt(c.dt, ct, t, rowmap_seq(2 * l, 0, 0, 0, l), nb, false, e.s);
                ct = t;
            }
            launch_mul_plain(c.dt, ct, pt, dst, 2, l, i > 0, nb, e.s);
        }
        if (bfv) {
            launch_ntt(c.dt, acc, acc, rowmap_seq(2 * l, 0, 0, 0, l), nb, true, e.s);
            if (has_acc) launch_binary(c.dt, OP_ADD, acc, adv(acc_in, b0), adv(out, b0), rowmap_seq(2 * l, 0, 0, 0, l), 2 * l, nb, e.s);
            else launch_binary(c.dt, OP_ADD, acc, acc, adv(out, b0), rowmap_seq(2 * l, 0, 0, 0, l), 0, nb, e.s);
        } else if (has_acc) {
            launch_binary(c.dt, OP_ADD, acc, adv(acc_in, b0), adv(out, b0), rowmap_seq(2 * l, 0, 0, 0, l), 2 * l, nb, e.s);
        }
    }
}

// out_b = sum_{t<n_mac} ct[t][b] * pt[t][b] + sum_{u<n_plain} ct[n_mac+u][b]   (tables are device arrays [term][batch]).
// Products need NTT-domain operands, i.e. CKKS ciphertexts and plaintexts; plain terms are limb-wise in any domain.
void op_sum_terms(OpEnv& e, int level, int npoly, const u64* const* ct_tab, const u64* const* pt_tab, u64* const* out_tab,
                  int n_mac, int n_plain, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    LSB_CHECK(level >= 0 && level < e.c.nq && npoly >= 1 && n_mac >= 0 && n_plain >= 0 && n_mac + n_plain >= 1, "lsb: sum_terms: bad arguments");
    LSB_CHECK(n_mac == 0 || e.c.scheme == SCHEME_CKKS, "lsb: sum_terms: fused products need NTT-domain (CKKS) operands");
    const int split = sum_terms_split(e.c.dt, npoly, level + 1, n_mac + n_plain, batch);
    u64* partial = split > 1 ? e.ws.reserve((size_t)split * batch * npoly * (level + 1) * e.c.N, e.s) : nullptr;
    launch_sum_terms(e.c.dt, ct_tab, pt_tab, out_tab, npoly, level + 1, n_mac, n_plain, batch, e.s, split, partial);
}

// ---------------------------------------------------------------------------------------------
// fused chains
// ---------------------------------------------------------------------------------------------
// Two-stream batch pipeline (LSB_DUAL_STREAM, default on for CKKS batches of >= 256): the batch is cut in two halves
// that run the same kernel chain on two streams, the second half one kernel behind the first, so that an FP64-pipe-bound
// NTT launch of one half shares the SMs with an HBM-bound limb-wise launch of the other and the tail of every launch is
// filled by the other stream's CTAs.  Same kernels on the same data: limbs are unchanged.
static int dual_stream_min_batch() {
    static const int v = [] {
        const char* e = getenv("LSB_DUAL_STREAM");
        if (e && atoi(e) == 0) return 1 << 30;
        return e && atoi(e) > 1 ? atoi(e) : 256;
    }();
    return v;
}
static void mult_relin_chain(OpEnv& e, int level, BatchPtr a, BatchPtr b, const u64* rlk, int key_level, BatchPtr out2,
                             int batch, bool rescale) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    Context& c = e.c;
    const int l = level + 1;
    const size_t N = c.N;
    const int mult_limbs = c.scheme == SCHEME_BFV ? bfv_mult_scratch_limbs(c, level) : 0;
    const int rest = std::max(relin_scratch_limbs(c, level), rescale ? rescale_scratch_limbs(c, level, 2) : 0);
    // [ct3 | ct2 (only when rescaling) | max(mult, relin, rescale) scratch]
    const int head = 3 * l + (rescale ? 2 * l : 0);
    const size_t item = (size_t)(head + std::max(mult_limbs, rest)) * N;
    auto run_chunk = [&](OpEnv& env, u64* scratch, int b0, int nb, cudaEvent_t after_first) {
        BatchPtr ct3 = bp_strided(scratch, item);
        BatchPtr ct2 = rescale ? bp_strided(scratch + (size_t)3 * l * N, item) : adv(out2, b0);
        u64* sub = scratch + (size_t)head * N;
        if (c.scheme == SCHEME_CKKS) ckks_tensor(env, level, adv(a, b0), adv(b, b0), ct3, nb);
        else bfv_mult_impl(env, level, adv(a, b0), adv(b, b0), ct3, sub, item, nb);
        if (after_first) LSB_CUDA(cudaEventRecord(after_first, env.s));
        relin_impl(env, level, ct3, rlk, key_level, ct2, sub, item, nb);
        if (rescale) rescale_impl(env, level, 2, ct2, adv(out2, b0), sub, item, nb);
    };
    const bool profiling = c.dt.prof && c.dt.prof->enabled.load(std::memory_order_relaxed);
    if (c.scheme == SCHEME_CKKS && e.chunk <= 0 && batch >= dual_stream_min_batch() && !profiling &&
        item * (size_t)batch <= scratch_budget_words()) {
        const int h0 = (batch + 1) / 2, h1 = batch - h0;
        u64* scratch = e.ws.reserve(item * batch, e.s);
        Context::SidePipe& sp = c.side_pipe();
        OpEnv side{e.c, e.ws, sp.stream, e.chunk};
        // the side stream may only start once everything queued on the caller's stream (its inputs) is done, and then
        // one kernel behind the first half
        run_chunk(e, scratch, 0, h0, sp.fork);
        LSB_CUDA(cudaStreamWaitEvent(sp.stream, sp.fork, 0));
        run_chunk(side, scratch + item * h0, h0, h1, nullptr);
        LSB_CUDA(cudaEventRecord(sp.join, sp.stream));
        LSB_CUDA(cudaStreamWaitEvent(e.s, sp.join, 0));
        return;
    }
    const int chunk = pick_chunk(e, item, batch, c.beta(level) * (l + c.np) - l);
    u64* scratch = e.ws.reserve(item * chunk, e.s);
    for (int b0 = 0; b0 < batch; b0 += chunk) run_chunk(e, scratch, b0, std::min(chunk, batch - b0), nullptr);
}
void op_mult_relin(OpEnv& e, int level, BatchPtr a, BatchPtr b, const u64* rlk, int key_level, BatchPtr out2,
                   int batch) {
    mult_relin_chain(e, level, a, b, rlk, key_level, out2, batch, false);
}
void op_mult_relin_rescale(OpEnv& e, int level, BatchPtr a, BatchPtr b, const u64* rlk, int key_level, BatchPtr out2,
                           int batch) {
    mult_relin_chain(e, level, a, b, rlk, key_level, out2, batch, true);
}

void op_ntt(OpEnv& e, BatchPtr in, BatchPtr out, int rows, const int* mod_idx, int period, bool inverse, int batch) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    LSB_CHECK(rows <= kMaxRows && period >= 1, "lsb: op_ntt rows/period out of range");
    RowMap m;
    m.n = rows;
    for (int r = 0; r < rows; r++) {
        m.in_slot[r] = m.out_slot[r] = (uint8_t)r;
        LSB_CHECK(mod_idx[r % period] >= 0 && mod_idx[r % period] < e.c.dt.nmod, "lsb: modulus index out of range");
        m.mod[r] = (uint8_t)mod_idx[r % period];
    }
    launch_ntt(e.c.dt, in, out, m, batch, inverse, e.s);
}

}  // namespace lsb
