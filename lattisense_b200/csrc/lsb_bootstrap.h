// CKKS bootstrapping as a pipeline of ciphertext operations — host-side orchestration only.
//
// Reference: the `bootstrap` node, mega_ag_runners/gpu/mega_ag_executors_gpu.cu:410-426 (HEonGPU
// regular_bootstrapping_v2(in, glk, rlk, &swk_dts, &swk_std)), its parameters mega_ag_runners/gpu/gpu_wrapper.cu:106-116
// and frontend/custom_task.py:424-463 (CtS depth 4 from level 24, EvalMod Cos1 K=16 degree 30 with 3 double-angle
// steps from level 20, StC depth 3 from level 12, output level 9, sparse-secret encapsulation keys swk_dts / swk_std
// :1989-1996), CPU twin mega_ag_runners/cpu/mega_ag_executors_cpu.cpp:451-468, acceptance unittests/test_gpu_ckks.cpp:743-808
// (decrypt + decode, mean precision >= 10 bits).  Both reference back ends delegate the algorithm to libraries that are
// absent from the tree (Lattigo / HEonGPU); what is pinned by in-tree sources is the op inventory, the level budget, the
// key set (frontend/bootstrap_params.py:96-263 fixes the rotations each CoeffsToSlots / SlotsToCoeffs matrix may use) and
// the precision bound.  This file follows those:
//
//   ScaleUp (x round(q0 / (message_ratio * scale)))  ->  key switch to the sparse secret (swk_dts, level 0)
//   -> ModRaise to the CtS start level -> key switch back to the dense secret (swk_std)
//   -> CoeffsToSlots: the special FFT of the canonical embedding factored into butterfly stages, merged into `depth`
//      diagonal matrices (same merge plan, same rotation sets and the same baby-step / giant-step split as the frontend),
//      each one level; real / imaginary parts split with ONE conjugation and an exact multiplication by the monomial
//      -X^(N/2) (slot-wise -i; no level)
//   -> EvalMod on both halves: Chebyshev interpolant of cos(2 pi (K u - 1/4) / 2^r) on [-1, 1] evaluated as a binary
//      tree over T_1, T_2, T_4, T_8, T_16 (depth ceil(log2(deg + 1)) exactly: every leaf is c0 + c1 T_1, whose constant
//      multiplication uses the level T_1 still has to spare), target scales set top-down so that every addition meets
//      equal scales; then r double-angle steps
//   -> recombine y_re + i y_im (monomial X^(N/2)) -> SlotsToCoeffs (depth matrices) with the factor that returns the
//      message to its input scale folded into the diagonals.
//
// The code is a template over the ciphertext operations (`Ops`): the product instantiates it with the device operations
// (lsb_bootstrap_dev.cu), the parity test with the CPU oracle's (tests/btp/btp_oracle.cpp), so the two runs execute the
// SAME sequence on the SAME encoded plaintexts and can be compared limb by limb.  Nothing here touches CUDA or oracle/.
#pragma once
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdint>
#include <map>
#include <memory>
#include <set>
#include <stdexcept>
#include <string>
#include <vector>

namespace lsb {
namespace btp {

typedef uint64_t u64;
typedef __int128 i128;
typedef std::complex<double> cplx;
typedef std::map<int, std::vector<cplx>> DiagMatrix;  // rotation k -> diagonal d_k:  (M v)[j] = sum_k d_k[j] v[(j + k) mod n]

struct Params {
    int logN = 0;
    int log_slots = 0;
    std::vector<u64> q, p;
    double scale = 0;
    int cts_start = 0, cts_depth = 0;
    double cts_ratio = 2.0;
    int evalmod_start = 0;
    double message_ratio = 256.0;
    int K = 16, sine_deg = 30, double_angle = 3;
    int stc_start = 0, stc_depth = 0;
    double stc_ratio = 2.0;
    int out_level = 0;
};

static const double kPi = 3.14159265358979323846264338327950288;

// ---- the special FFT's factorisation (complex model; tests/btp_model.py is the numpy twin) -----------------------
inline std::vector<u64> rot_group(int n) {  // 5^j mod 4n
    std::vector<u64> out(n);
    u64 g = 1;
    for (int j = 0; j < n; j++) {
        out[j] = g;
        g = g * 5 % (4ull * n);
    }
    return out;
}
inline DiagMatrix butterfly_stage(int n, int length, bool c2s) {
    const int h = length / 2;
    const std::vector<u64> rg = rot_group(n);
    std::vector<cplx> d0(n), dp(n), dm(n);
    for (int x = 0; x < n; x++) {
        const int j = x % length;
        const double ang = 2.0 * kPi * (double)((rg[j % h] % (4ull * length)) * (u64)(n / length)) / (4.0 * n);
        const bool lo = j < h;
        if (!c2s) {
            const cplx tw(std::cos(ang), std::sin(ang));
            d0[x] = lo ? cplx(1.0, 0.0) : -tw;
            dp[x] = lo ? tw : cplx(0.0, 0.0);
            dm[x] = lo ? cplx(0.0, 0.0) : cplx(1.0, 0.0);
        } else {
            const cplx twi(std::cos(ang), -std::sin(ang));
            d0[x] = lo ? cplx(0.5, 0.0) : -0.5 * twi;
            dp[x] = lo ? cplx(0.5, 0.0) : cplx(0.0, 0.0);
            dm[x] = lo ? cplx(0.0, 0.0) : 0.5 * twi;
        }
    }
    DiagMatrix m;
    m[0] = d0;
    if (h % n != (n - h) % n) {
        m[h % n] = dp;
        m[(n - h) % n] = dm;
    } else {
        for (int x = 0; x < n; x++) dp[x] += dm[x];
        m[h % n] = dp;
    }
    return m;
}
// diagonals of A*B (B applied first): (AB)_{i+j} += A_i * rot(B_j, i)
inline DiagMatrix compose(const DiagMatrix& a, const DiagMatrix& b) {
    const int n = (int)a.begin()->second.size();
    DiagMatrix out;
    for (const auto& ia : a)
        for (const auto& jb : b) {
            const int k = (ia.first + jb.first) % n;
            std::vector<cplx>& o = out[k];
            if (o.empty()) o.assign(n, cplx(0.0, 0.0));
            for (int x = 0; x < n; x++) o[x] += ia.second[x] * jb.second[(x + ia.first) % n];
        }
    return out;
}
// how many butterfly stages each of the `depth` matrices absorbs (frontend/bootstrap_params.py:104-112)
inline std::vector<int> merge_plan(int log_slots, int depth, bool c2s) {
    std::vector<int> merge(depth, 0);
    int level = log_slots;
    for (int i = 0; i < depth; i++) {
        const int d = (level + (depth - i) - 1) / (depth - i);
        merge[c2s ? i : depth - i - 1] = d;
        level -= d;
    }
    return merge;
}
inline std::vector<DiagMatrix> factor_c2s(int n, int depth) {  // applied in order; result = BitRev(U^-1 z)
    int log_n = 0;
    while ((1 << log_n) < n) log_n++;
    std::vector<DiagMatrix> mats;
    int level = log_n;
    for (int m : merge_plan(log_n, depth, true)) {
        DiagMatrix cur = butterfly_stage(n, 1 << level, true);
        for (int s = 1; s < m; s++) cur = compose(butterfly_stage(n, 1 << (level - s), true), cur);
        mats.push_back(std::move(cur));
        level -= m;
    }
    return mats;
}
inline std::vector<DiagMatrix> factor_s2c(int n, int depth) {  // applied in order to BitRev(w); result = U w
    int log_n = 0;
    while ((1 << log_n) < n) log_n++;
    std::vector<DiagMatrix> mats;
    int level = log_n;
    for (int m : merge_plan(log_n, depth, false)) {
        DiagMatrix cur = butterfly_stage(n, 2 << (log_n - level), false);
        for (int s = 1; s < m; s++) cur = compose(butterfly_stage(n, 2 << (log_n - level + s), false), cur);
        mats.push_back(std::move(cur));
        level -= m;
    }
    return mats;
}
// rotation set of each merged matrix (frontend/bootstrap_params.py:96-139, full-slot case)
inline std::vector<std::vector<int>> dft_index_map(int log_slots, int depth, bool c2s) {
    const int n = 1 << log_slots;
    std::vector<std::vector<int>> out;
    int level = log_slots;
    for (int m : merge_plan(log_slots, depth, c2s)) {
        std::set<int> rots{0};
        for (int s = 0; s < m; s++) {
            const int lv = level - s;
            const int r = c2s ? (1 << (lv - 1)) : (1 << (log_slots - lv));
            std::set<int> next;
            for (int a : rots)
                for (int b : {0, r, n - r}) next.insert((a + b) % n);
            rots.swap(next);
        }
        out.emplace_back(rots.begin(), rots.end());
        level -= m;
    }
    return out;
}
// n1 such that #baby / #giant rotations is closest below `ratio` (frontend/bootstrap_params.py:188-201)
inline int bsgs_split(const std::vector<int>& rots, int n, double ratio) {
    auto counts = [&](int n1, int& giant, int& baby) {
        std::set<int> g, b;
        for (int r : rots) {
            g.insert(((r / n1) * n1) % n);
            b.insert(r % n1);
        }
        giant = (int)g.size() - 1;
        baby = (int)b.size() - 1;
    };
    for (int n1 = 1; n1 < n; n1 <<= 1) {
        int g, b;
        counts(n1, g, b);
        if (g && (double)b / g == ratio) return n1;
        if (g == 0 || (double)b / g > ratio) return std::max(n1 / 2, 1);
    }
    return 1;
}

// ---- canonical-embedding encoder (slot j <-> evaluation at xi^(5^j), xi = exp(i pi / N)) ---------------------------
// O(n log n) inverse special FFT (decimation in frequency over the rotation group, then bit reversal), full packing:
// n = N/2 slots, coefficient k = Re w_k, coefficient k + n = Im w_k.
struct Encoder {
    int N, n;
    std::vector<u64> rg;       // 5^j mod 2N
    std::vector<cplx> ksi;     // exp(2 pi i k / 2N), k <= 2N
    explicit Encoder(int N_) : N(N_), n(N_ / 2), rg(N_ / 2), ksi(2 * (size_t)N_ + 1) {
        u64 g = 1;
        for (int j = 0; j < n; j++) {
            rg[j] = g;
            g = g * 5 % (2ull * N);
        }
        for (size_t k = 0; k <= 2 * (size_t)N; k++) {
            const double a = 2.0 * kPi * (double)k / (2.0 * N);
            ksi[k] = cplx(std::cos(a), std::sin(a));
        }
    }
    // slots (n values) -> packed coefficients w (n values)
    void embed_inv(std::vector<cplx>& v) const {
        const long M = 2L * N;
        for (long len = n; len >= 1; len >>= 1) {
            const long lenh = len >> 1, lenq = len << 2, gap = M / lenq;
            for (long i = 0; i < n; i += len)
                for (long j = 0; j < lenh; j++) {
                    const long idx = (lenq - (long)(rg[j] % (u64)lenq)) * gap;
                    const cplx u = v[i + j] + v[i + j + lenh];
                    const cplx w = (v[i + j] - v[i + j + lenh]) * ksi[idx];
                    v[i + j] = u;
                    v[i + j + lenh] = w;
                }
        }
        int bits = 0;
        while ((1 << bits) < n) bits++;
        for (int i = 0; i < n; i++) {
            int r = 0;
            for (int b = 0; b < bits; b++) r |= ((i >> b) & 1) << (bits - 1 - b);
            if (i < r) std::swap(v[i], v[r]);
        }
        const double inv = 1.0 / n;
        for (int i = 0; i < n; i++) v[i] *= inv;
    }
    // round(scale * coefficient) reduced modulo q_0..q_level: coefficient-domain limbs [level+1][N]
    std::vector<u64> encode(std::vector<cplx> slots, int level, long double scale, const std::vector<u64>& q) const {
        if ((int)slots.size() != n) throw std::runtime_error("lsb bootstrap: encode expects N/2 slot values");
        embed_inv(slots);
        std::vector<u64> out((size_t)(level + 1) * N);
        for (int k = 0; k < N; k++) {
            const long double c = (long double)(k < n ? slots[k].real() : slots[k - n].imag()) * scale;
            if (!(fabsl(c) < 1.0e37L)) throw std::runtime_error("lsb bootstrap: plaintext coefficient out of range");
            const i128 v = (i128)roundl(c);
            for (int i = 0; i <= level; i++) {
                const i128 m = (i128)q[i];
                i128 x = v % m;
                if (x < 0) x += m;
                out[(size_t)i * N + k] = (u64)x;
            }
        }
        return out;
    }
};

// integer (any sign) -> residues modulo q_0..q_level
inline std::vector<u64> residues(i128 v, int level, const std::vector<u64>& q) {
    std::vector<u64> out(level + 1);
    for (int i = 0; i <= level; i++) {
        const i128 m = (i128)q[i];
        i128 x = v % m;
        if (x < 0) x += m;
        out[i] = (u64)x;
    }
    return out;
}
inline i128 round_i128(long double v) { return (i128)roundl(v); }

// Chebyshev interpolation coefficients of f on [-1, 1] (nodes of T_{deg+1})
template <class F>
inline std::vector<long double> chebyshev_coeffs(F&& f, int deg) {
    const int m = deg + 1;
    const long double pi = 3.14159265358979323846264338327950288L;
    std::vector<long double> y(m), c(m);
    for (int k = 0; k < m; k++) y[k] = f(cosl(pi * (k + 0.5L) / m));
    for (int j = 0; j < m; j++) {
        long double s = 0;
        for (int k = 0; k < m; k++) s += y[k] * cosl(pi * j * (k + 0.5L) / m);
        c[j] = 2.0L / m * s;
    }
    c[0] *= 0.5L;
    return c;
}
// p = lo + T_m * hi for a Chebyshev series of 2m coefficients (T_{m+i} = 2 T_m T_i - T_{m-i})
inline void chebyshev_split(const std::vector<long double>& c, int m, std::vector<long double>& lo, std::vector<long double>& hi) {
    lo.assign(c.begin(), c.begin() + m);
    hi.assign(m, 0.0L);
    hi[0] = c[m];
    for (int i = 1; i < m; i++) {
        hi[i] = 2.0L * c[m + i];
        lo[m - i] -= c[m + i];
    }
}

// ---- the pipeline ---------------------------------------------------------------------------------------------------
// Ops concept (batched: a Ct is a batch of ciphertexts of one level, a Pt one encoded plaintext shared by the batch):
//   typedef Ct, Pt (movable handles);  int level(const Ct&)
//   Pt make_plain(int level, const std::vector<u64>& coeff_limbs)   coefficient-domain residues [level+1][N] -> NTT domain
//   Ct rotate(const Ct&, int step)    Ct conjugate(const Ct&)        Galois keys 5^step mod 2N / 2N-1
//   vector<Ct> rotate_many(const Ct&, steps)                          the same rotations, free to share the decomposition
//   Ct switch_key(const Ct&, int which)                               0: swk_dts, 1: swk_std (no automorphism)
//   Ct mod_raise(const Ct& level0, int top)
//   Ct add(a, b)  Ct sub(a, b)  Ct mul_relin(a, b)  Ct rescale(a)  Ct drop_to(a, level)
//   Ct mul_scalar(a, residues)   both polynomials times an integer constant given modulo q_0..q_level
//   Ct add_scalar(a, residues)   constant added to every NTT slot of c0 (a constant polynomial)
//   Ct mul_plain(a, pt)          Ct sum_products(cts, pts) = sum_i ct_i * pt_i      Ct sum(cts)
//   Ct concat(a, b): the items of a followed by the items of b as one batch;   pair<Ct, Ct> split(c): its inverse
template <class Ops>
class Bootstrapper {
public:
    typedef typename Ops::Ct Ct;
    typedef typename Ops::Pt Pt;
    struct Sc {  // ciphertext + the scale its plaintext is read at
        Ct ct;
        long double scale;
    };
    struct Stats {
        int rotations = 0, key_switches = 0, mults = 0, rescales = 0, plain_mults = 0;
    };

    Bootstrapper(const Params& prm) : P(prm), N(1 << prm.logN), n(1 << prm.log_slots), enc(1 << prm.logN) {
        if (n != N / 2) throw std::runtime_error("lsb bootstrap: sparse packing (slots < N/2) is not supported by this backend");
        if (P.cts_start - P.cts_depth != P.evalmod_start) throw std::runtime_error("lsb bootstrap: CoeffsToSlots must end at the EvalMod start level");
        int k = 0;
        while ((1 << k) < P.sine_deg + 1) k++;
        tree_depth = k;
        if (P.evalmod_start - tree_depth - P.double_angle != P.stc_start) throw std::runtime_error("lsb bootstrap: EvalMod depth does not meet the SlotsToCoeffs start level");
        if (P.stc_start - P.stc_depth != P.out_level) throw std::runtime_error("lsb bootstrap: SlotsToCoeffs does not end at the output level");
        if (P.cts_start > (int)P.q.size() - 1) throw std::runtime_error("lsb bootstrap: start level outside the modulus chain");
        cts_mats = factor_c2s(n, P.cts_depth);
        stc_mats = factor_s2c(n, P.stc_depth);
        const auto ci = dft_index_map(P.log_slots, P.cts_depth, true), si = dft_index_map(P.log_slots, P.stc_depth, false);
        for (int i = 0; i < P.cts_depth; i++) cts_n1.push_back(bsgs_split(ci[i], n, P.cts_ratio));
        for (int i = 0; i < P.stc_depth; i++) stc_n1.push_back(bsgs_split(si[i], n, P.stc_ratio));
        // ScaleUp constant: the message goes to q0 / message_ratio before the modulus is raised
        scale_up = (i128)roundl((long double)P.q[0] / ((long double)P.message_ratio * (long double)P.scale));
        if (scale_up < 1) scale_up = 1;
        const int r = P.double_angle, K = P.K;
        cheb = chebyshev_coeffs(
            [&](long double u) { return cosl(2.0L * 3.14159265358979323846264338327950288L * ((long double)K * u - 0.25L) / (long double)(1 << r)); },
            P.sine_deg);
        cheb.resize((size_t)1 << tree_depth, 0.0L);
    }

    // rotation steps (mod N/2) the pipeline asks Ops::rotate for -- a subset of the frontend's rotations_for_bootstrapping()
    std::vector<int> rotations() const {
        std::set<int> out;
        auto collect = [&](const std::vector<DiagMatrix>& mats, const std::vector<int>& n1s) {
            for (size_t i = 0; i < mats.size(); i++)
                for (const auto& kd : mats[i]) {
                    const int k = kd.first, n1 = n1s[i];
                    if (k % n1 != 0) out.insert(k % n1);
                    if (k / n1 != 0) out.insert((k / n1) * n1);
                }
        };
        collect(cts_mats, cts_n1);
        collect(stc_mats, stc_n1);
        return std::vector<int>(out.begin(), out.end());
    }

    // Encode every plaintext the pipeline multiplies by (done once per Ops instance; levels and factors are fixed by the
    // parameters, only S_e -- the scale EvalMod ends at -- has to be derived by a dry run of the scale arithmetic).
    void precompute(Ops& ops) {
        if (ready) return;
        const long double cts_factor = powl(1.0L / (2.0L * P.K), 1.0L / P.cts_depth);
        for (int i = 0; i < P.cts_depth; i++) cts_plain.push_back(encode_matrix(ops, cts_mats[i], cts_n1[i], P.cts_start - i, cts_factor));
        const long double S1 = (long double)P.q[0];
        const long double Se = evalmod_scale(S1);
        // sin(2 pi x) ~ 2 pi * scale_up * P_k / q0 at scale Se  ->  P_k at the input scale
        const long double F = (long double)P.q[0] / (2.0L * 3.14159265358979323846264338327950288L * (long double)scale_up * Se);
        const long double stc_factor = powl(F, 1.0L / P.stc_depth);
        for (int i = 0; i < P.stc_depth; i++) stc_plain.push_back(encode_matrix(ops, stc_mats[i], stc_n1[i], P.stc_start - i, stc_factor));
        mono_neg_i = ops.make_plain(P.evalmod_start, monomial(P.evalmod_start, true));
        mono_i = ops.make_plain(P.stc_start, monomial(P.stc_start, false));
        ready = true;
    }

    Ct run(Ops& ops, const Ct& ct0) {
        precompute(ops);
        if (ops.level(ct0) != 0) throw std::runtime_error("lsb bootstrap: input ciphertext must be at level 0");
        Ct c = ops.mul_scalar(ct0, residues(scale_up, 0, P.q));
        c = ops.switch_key(c, 0);
        Ct r = ops.mod_raise(c, P.cts_start);
        r = ops.switch_key(r, 1);
        st.key_switches += 2;
        Sc v{std::move(r), (long double)P.q[0]};  // plaintext read in units of q0: x = scale_up * m * Delta / q0 + I
        for (int i = 0; i < P.cts_depth; i++) v = linear(ops, v, cts_plain[i]);
        Ct vc = ops.conjugate(v.ct);
        st.key_switches++;
        // real and imaginary halves go through EvalMod as ONE stacked batch (the two evaluations are the same op sequence
        // on independent ciphertexts): half the launches, twice the work per launch
        Sc t{ops.concat(ops.add(v.ct, vc), ops.mul_plain(ops.sub(v.ct, vc), mono_neg_i)), v.scale};
        Sc y = eval_mod(ops, t);
        std::pair<Ct, Ct> halves = ops.split(y.ct);
        Sc z{ops.add(halves.first, ops.mul_plain(halves.second, mono_i)), y.scale};
        for (int i = 0; i < P.stc_depth; i++) z = linear(ops, z, stc_plain[i]);
        return std::move(z.ct);
    }

    Stats st;
    const Params P;

private:
    struct EncodedMatrix {
        int level, n1;
        long double pt_scale;
        std::vector<int> rot;  // diagonal index k of each plaintext
        std::vector<Pt> pt;    // diagonal k rotated by -(giant step of k), times the matrix' share of the constant factor
    };
    const int N, n;
    Encoder enc;
    int tree_depth = 0;
    i128 scale_up = 1;
    std::vector<long double> cheb;
    std::vector<DiagMatrix> cts_mats, stc_mats;
    std::vector<int> cts_n1, stc_n1;
    std::vector<EncodedMatrix> cts_plain, stc_plain;
    Pt mono_i, mono_neg_i;
    bool ready = false;

    std::vector<u64> monomial(int level, bool negative) const {  // +-X^(N/2): slot-wise multiplication by +-i
        std::vector<u64> m((size_t)(level + 1) * N, 0);
        for (int i = 0; i <= level; i++) m[(size_t)i * N + N / 2] = negative ? P.q[i] - 1 : 1;
        return m;
    }
    EncodedMatrix encode_matrix(Ops& ops, const DiagMatrix& M, int n1, int level, long double factor) {
        EncodedMatrix e;
        e.level = level;
        e.n1 = n1;
        e.pt_scale = (long double)P.q[level];
        for (const auto& kd : M) {
            const int k = kd.first, g = (k / n1) * n1;
            std::vector<cplx> d(n);
            for (int x = 0; x < n; x++) d[x] = kd.second[((x - g) % n + n) % n] * (double)factor;  // np.roll(d, g)
            e.rot.push_back(k);
            e.pt.push_back(ops.make_plain(level, enc.encode(std::move(d), level, e.pt_scale, P.q)));
        }
        return e;
    }
    // sum over giant steps g of rot( sum_b pt_{g+b} * rot(ct, b), g ), then one rescale by q_level (= the plaintext scale)
    Sc linear(Ops& ops, const Sc& in, const EncodedMatrix& e) {
        if (ops.level(in.ct) != e.level) throw std::runtime_error("lsb bootstrap: linear transform at the wrong level");
        std::map<int, Ct> baby;
        std::map<int, std::vector<std::pair<int, const Pt*>>> groups;
        std::vector<int> steps;
        for (size_t i = 0; i < e.rot.size(); i++) {
            const int k = e.rot[i], b = k % e.n1, g = k - b;
            groups[g].push_back({b, &e.pt[i]});
            if (b && std::find(steps.begin(), steps.end(), b) == steps.end()) steps.push_back(b);
        }
        if (!steps.empty()) {
            // all baby steps rotate the same ciphertext: one shared decomposition (hoisting), limbs as rotate()'s
            std::vector<Ct> rot = ops.rotate_many(in.ct, steps);
            for (size_t i = 0; i < steps.size(); i++) baby.emplace(steps[i], std::move(rot[i]));
            st.rotations += (int)steps.size();
        }
        std::vector<Ct> parts;
        for (auto& gi : groups) {
            std::vector<const Ct*> cts;
            std::vector<const Pt*> pts;
            for (auto& bp : gi.second) {
                cts.push_back(bp.first ? &baby.at(bp.first) : &in.ct);
                pts.push_back(bp.second);
            }
            st.plain_mults += (int)cts.size();
            Ct acc = ops.sum_products(cts, pts);
            if (gi.first) {
                acc = ops.rotate(acc, gi.first);
                st.rotations++;
            }
            parts.push_back(std::move(acc));
        }
        std::vector<const Ct*> pp;
        for (auto& c : parts) pp.push_back(&c);
        Ct total = parts.size() == 1 ? std::move(parts[0]) : ops.sum(pp);
        st.rescales++;
        return Sc{ops.rescale(total), in.scale * e.pt_scale / (long double)P.q[e.level]};
    }

    // ---- EvalMod ----
    struct Basis {
        std::vector<Sc> T;  // T[j] = T_{2^j}(u), level evalmod_start - j
    };
    long double q_at(int level) const { return (long double)P.q[level]; }
    // scale arithmetic of dbl() without touching ciphertexts
    static long double dbl_scale(long double s, long double ql) { return s * s / ql; }
    long double evalmod_scale(long double S1) const {
        // the tree returns its result at the scale it is asked for; we ask for T_1's scale, then r doublings
        long double s = S1;
        int level = P.evalmod_start - tree_depth;
        for (int i = 0; i < P.double_angle; i++) {
            s = dbl_scale(s, q_at(level));
            level--;
        }
        return s;
    }
    // 2 a^2 - 1
    Sc dbl(Ops& ops, const Sc& a) {
        const int level = ops.level(a.ct);
        Ct sq = ops.mul_relin(a.ct, a.ct);
        sq = ops.mul_scalar(sq, residues(2, level, P.q));
        Ct r = ops.rescale(sq);
        st.mults++;
        st.key_switches++;
        st.rescales++;
        const long double s = dbl_scale(a.scale, q_at(level));
        return Sc{ops.add_scalar(r, residues(-round_i128(s), level - 1, P.q)), s};
    }
    // sum_{j < 2^k} c_j T_j at level evalmod_start - k, plaintext scale `scale`
    Sc tree(Ops& ops, const Basis& B, const std::vector<long double>& c, int k, long double scale) {
        const int top = P.evalmod_start;
        if (k == 1) {
            const Sc& T1 = B.T[0];
            const long double cs = scale * q_at(top) / T1.scale;
            Ct m = ops.rescale(ops.mul_scalar(T1.ct, residues(round_i128(c[1] * cs), top, P.q)));
            st.rescales++;
            return Sc{ops.add_scalar(m, residues(round_i128(c[0] * scale), top - 1, P.q)), scale};
        }
        const int m = 1 << (k - 1);
        std::vector<long double> lo, hi;
        chebyshev_split(c, m, lo, hi);
        const Sc& Tm = B.T[k - 1];
        const int lv = top - (k - 1);  // level of T_m and of both halves
        Sc L = tree(ops, B, lo, k - 1, scale);
        bool hi_zero = true;
        for (long double h : hi) hi_zero = hi_zero && fabsl(h) < 1e-30L;
        if (hi_zero) return Sc{ops.drop_to(L.ct, lv - 1), scale};
        Sc H = tree(ops, B, hi, k - 1, scale * q_at(lv) / Tm.scale);
        Ct prod = ops.rescale(ops.mul_relin(H.ct, Tm.ct));
        st.mults++;
        st.key_switches++;
        st.rescales++;
        return Sc{ops.add(prod, ops.drop_to(L.ct, lv - 1)), scale};
    }
    Sc eval_mod(Ops& ops, const Sc& t) {
        Basis B;
        B.T.push_back(Sc{ops.drop_to(t.ct, P.evalmod_start), t.scale});
        for (int j = 1; j < tree_depth; j++) B.T.push_back(dbl(ops, B.T[j - 1]));
        Sc y = tree(ops, B, cheb, tree_depth, t.scale);
        for (int i = 0; i < P.double_angle; i++) y = dbl(ops, y);
        return y;
    }
};

}  // namespace btp
}  // namespace lsb
