// TEST INFRASTRUCTURE: the product's bootstrapping pipeline (lattisense_b200/csrc/lsb_bootstrap.h, a template over the
// ciphertext operations) instantiated with the CPU ORACLE's operations.  The device run (lsb_bootstrap_dev.cu) and this
// run execute the same sequence on the same encoded plaintexts, so tests compare their output limbs bit for bit; the
// pipeline itself is validated by decrypt + decode precision (the reference's criterion, unittests/test_gpu_ckks.cpp:743-808).
// Never linked into the product.
#include <cstdio>
#include <cstring>
#include "../../lattisense_b200/csrc/lsb_bootstrap.h"
#include "../../oracle/orc_keys.hpp"

using orc::u64;

struct OracleOps {
    struct Ct {
        int level = -1;
        int nb = 0;          // batch items held (the pipeline stacks two batches for EvalMod)
        std::vector<u64> d;  // [nb][2][level+1][N]
    };
    struct Pt {
        int level = -1;
        std::vector<u64> d;  // [level+1][N], NTT domain
    };
    const orc::Context& c;
    int batch;
    int key_level;
    const u64* rlk;
    std::map<u64, const u64*> glk;
    const u64* swk[2];
    int swk_level[2];
    size_t N;

    OracleOps(const orc::Context& ctx, int batch_) : c(ctx), batch(batch_), N(ctx.N) {}
    int level(const Ct& a) const { return a.level; }
    size_t words(int level) const { return (size_t)2 * (level + 1) * N; }
    Ct make(int level, int nb) const {
        Ct o;
        o.level = level;
        o.nb = nb;
        o.d.assign(words(level) * nb, 0);
        return o;
    }
    Ct concat(const Ct& a, const Ct& b) const {
        Ct o = make(a.level, a.nb + b.nb);
        std::memcpy(o.d.data(), a.d.data(), a.d.size() * sizeof(u64));
        std::memcpy(o.d.data() + a.d.size(), b.d.data(), b.d.size() * sizeof(u64));
        return o;
    }
    std::pair<Ct, Ct> split(const Ct& c) const {
        Ct a = make(c.level, c.nb / 2), b = make(c.level, c.nb - c.nb / 2);
        std::memcpy(a.d.data(), c.d.data(), a.d.size() * sizeof(u64));
        std::memcpy(b.d.data(), c.d.data() + a.d.size(), b.d.size() * sizeof(u64));
        return {std::move(a), std::move(b)};
    }
    Pt make_plain(int level, const std::vector<u64>& coeff) const {
        Pt p;
        p.level = level;
        p.d = coeff;
        for (int i = 0; i <= level; i++) orc::ntt_forward(c.Q(i), c.N, &p.d[(size_t)i * N]);
        return p;
    }
    u64 gal_of(int step) const {
        const u64 m = 2 * N;
        u64 r = 1, b = 5 % m;
        for (u64 e = (u64)step; e; e >>= 1, b = b * b % m)
            if (e & 1) r = r * b % m;
        return r;
    }
    Ct rotate_gal(const Ct& a, u64 gal) const {
        auto it = glk.find(gal);
        if (it == glk.end()) throw std::runtime_error("btp oracle: no Galois key for element " + std::to_string(gal));
        Ct o = make(a.level, a.nb);
        for (int b = 0; b < a.nb; b++)
            orc::ckks_rotate(c, a.level, &a.d[words(a.level) * b], it->second, key_level, gal, &o.d[words(a.level) * b]);
        return o;
    }
    Ct rotate(const Ct& a, int step) const { return rotate_gal(a, gal_of(step)); }
    Ct conjugate(const Ct& a) const { return rotate_gal(a, 2 * N - 1); }
    std::vector<Ct> rotate_many(const Ct& a, const std::vector<int>& steps) const {
        std::vector<Ct> out;
        for (int st : steps) out.push_back(rotate(a, st));
        return out;
    }
    Ct switch_key(const Ct& a, int which) const {
        const int l = a.level + 1;
        Ct o = make(a.level, a.nb);
        std::vector<u64> p0((size_t)l * N), p1((size_t)l * N);
        for (int b = 0; b < a.nb; b++) {
            const u64* ct = &a.d[words(a.level) * b];
            u64* out = &o.d[words(a.level) * b];
            orc::keyswitch_ntt(c, a.level, ct + (size_t)l * N, swk[which], swk_level[which], p0.data(), p1.data());
            for (int i = 0; i < l; i++) {
                const u64 q = c.Q(i).q;
                for (size_t x = 0; x < N; x++) {
                    out[(size_t)i * N + x] = orc::addmod(ct[(size_t)i * N + x], p0[(size_t)i * N + x], q);
                    out[(size_t)(l + i) * N + x] = p1[(size_t)i * N + x];
                }
            }
        }
        return o;
    }
    Ct mod_raise(const Ct& a, int top) const {
        if (a.level != 0) throw std::runtime_error("btp oracle: mod_raise expects level 0");
        Ct o = make(top, a.nb);
        const u64 q0 = c.Q(0).q;
        std::vector<u64> co(N);
        for (int b = 0; b < a.nb; b++)
            for (int p = 0; p < 2; p++) {
                std::memcpy(co.data(), &a.d[words(0) * b + (size_t)p * N], N * sizeof(u64));
                orc::ntt_inverse(c.Q(0), c.N, co.data());
                for (int i = 0; i <= top; i++) {
                    const u64 q = c.Q(i).q;
                    u64* dst = &o.d[words(top) * b + ((size_t)p * (top + 1) + i) * N];
                    for (size_t x = 0; x < N; x++) {
                        const u64 v = co[x];
                        dst[x] = v > q0 / 2 ? orc::negmod((q0 - v) % q, q) : v % q;
                    }
                    orc::ntt_forward(c.Q(i), c.N, dst);
                }
            }
        return o;
    }
    Ct add(const Ct& a, const Ct& b) const {
        Ct o = make(a.level, a.nb);
        for (int i = 0; i < a.nb; i++) orc::ct_add(c, a.level, 2, &a.d[words(a.level) * i], &b.d[words(a.level) * i], &o.d[words(a.level) * i]);
        return o;
    }
    Ct sub(const Ct& a, const Ct& b) const {
        Ct o = make(a.level, a.nb);
        for (int i = 0; i < a.nb; i++) orc::ct_sub(c, a.level, 2, &a.d[words(a.level) * i], &b.d[words(a.level) * i], &o.d[words(a.level) * i]);
        return o;
    }
    Ct mul_relin(const Ct& a, const Ct& b) const {
        Ct o = make(a.level, a.nb);
        std::vector<u64> t3((size_t)3 * (a.level + 1) * N);
        for (int i = 0; i < a.nb; i++) {
            orc::ckks_mult(c, a.level, &a.d[words(a.level) * i], &b.d[words(a.level) * i], t3.data());
            orc::ckks_relin(c, a.level, t3.data(), rlk, key_level, &o.d[words(a.level) * i]);
        }
        return o;
    }
    Ct rescale(const Ct& a) const {
        Ct o = make(a.level - 1, a.nb);
        for (int i = 0; i < a.nb; i++) orc::ckks_rescale(c, a.level, 2, &a.d[words(a.level) * i], &o.d[words(a.level - 1) * i]);
        return o;
    }
    Ct drop_to(const Ct& a, int level) const {
        if (level == a.level) return a;
        Ct o = make(level, a.nb);
        for (int i = 0; i < a.nb; i++)
            for (int p = 0; p < 2; p++)
                std::memcpy(&o.d[words(level) * i + (size_t)p * (level + 1) * N], &a.d[words(a.level) * i + (size_t)p * (a.level + 1) * N],
                            (size_t)(level + 1) * N * sizeof(u64));
        return o;
    }
    Ct mul_scalar(const Ct& a, const std::vector<u64>& k) const {
        Ct o = make(a.level, a.nb);
        const int l = a.level + 1;
        for (size_t r = 0; r < (size_t)a.nb * 2 * l; r++) {
            const int i = (int)(r % l);
            const u64 q = c.Q(i).q;
            for (size_t x = 0; x < N; x++) o.d[r * N + x] = orc::mulmod(a.d[r * N + x], k[i], q);
        }
        return o;
    }
    Ct add_scalar(const Ct& a, const std::vector<u64>& k) const {
        Ct o = a;
        const int l = a.level + 1;
        for (int b = 0; b < a.nb; b++)
            for (int i = 0; i < l; i++) {
                const u64 q = c.Q(i).q;
                u64* p = &o.d[words(a.level) * b + (size_t)i * N];
                for (size_t x = 0; x < N; x++) p[x] = orc::addmod(p[x], k[i], q);
            }
        return o;
    }
    Ct sum_products(const std::vector<const Ct*>& cts, const std::vector<const Pt*>& pts) const {
        const int lv = cts[0]->level, l = lv + 1;
        Ct o = make(lv, cts[0]->nb);
        for (size_t t = 0; t < cts.size(); t++)
            for (size_t r = 0; r < (size_t)cts[0]->nb * 2 * l; r++) {
                const int i = (int)(r % l);
                const u64 q = c.Q(i).q;
                const u64* pt = &pts[t]->d[(size_t)i * N];
                const u64* ct = &cts[t]->d[r * N];
                u64* out = &o.d[r * N];
                for (size_t x = 0; x < N; x++) out[x] = orc::addmod(out[x], orc::mulmod(ct[x], pt[x], q), q);
            }
        return o;
    }
    Ct mul_plain(const Ct& a, const Pt& p) const { return sum_products({&a}, {&p}); }
    Ct sum(const std::vector<const Ct*>& cts) const {
        Ct o = *cts[0];
        for (size_t t = 1; t < cts.size(); t++) o = add(o, *cts[t]);
        return o;
    }
};

extern "C" {

// params: [logN, log_slots, cts_start, cts_depth, evalmod_start, K, sine_deg, double_angle, stc_start, stc_depth, out_level]
// dparams: [scale, cts_ratio, message_ratio, stc_ratio]
// keys: gals[n_glk] / glk[n_glk] pointers (col rotations and the row element), all at key_level; swk_dts at level 0
// ct0: [batch][2][1][N];  out: [batch][2][out_level+1][N];  stats8: rotations, key_switches, mults, rescales, plain_mults
int btp_oracle_run(const int* ip, const double* dp, const u64* q, int nq, const u64* p, int np, int batch, const u64* ct0, const u64* rlk,
                   int key_level, const u64* gals, const u64* const* glk, int n_glk, const u64* swk_dts, const u64* swk_std, u64* out,
                   long long* stats8, int* rots_out, int rots_cap) {
    try {
        auto ctx = orc::make_context(orc::SCHEME_CKKS, ip[0], q, nq, p, np, 0);
        lsb::btp::Params P;
        P.logN = ip[0];
        P.log_slots = ip[1];
        P.q.assign(q, q + nq);
        P.p.assign(p, p + np);
        P.scale = dp[0];
        P.cts_start = ip[2];
        P.cts_depth = ip[3];
        P.cts_ratio = dp[1];
        P.evalmod_start = ip[4];
        P.message_ratio = dp[2];
        P.K = ip[5];
        P.sine_deg = ip[6];
        P.double_angle = ip[7];
        P.stc_start = ip[8];
        P.stc_depth = ip[9];
        P.stc_ratio = dp[3];
        P.out_level = ip[10];
        lsb::btp::Bootstrapper<OracleOps> B(P);
        if (rots_out) {
            auto r = B.rotations();
            int k = 0;
            for (int v : r)
                if (k < rots_cap) rots_out[k++] = v;
            if (k < rots_cap) rots_out[k] = -1;
        }
        if (!ct0) return 0;  // rotation query only
        OracleOps ops(*ctx, batch);
        ops.key_level = key_level;
        ops.rlk = rlk;
        for (int i = 0; i < n_glk; i++) ops.glk[gals[i]] = glk[i];
        ops.swk[0] = swk_dts;
        ops.swk[1] = swk_std;
        ops.swk_level[0] = 0;
        ops.swk_level[1] = key_level;
        OracleOps::Ct in = ops.make(0, batch);
        std::memcpy(in.d.data(), ct0, in.d.size() * sizeof(u64));
        OracleOps::Ct res = B.run(ops, in);
        if (res.level != P.out_level) throw std::runtime_error("btp oracle: unexpected output level");
        std::memcpy(out, res.d.data(), res.d.size() * sizeof(u64));
        if (stats8) {
            stats8[0] = B.st.rotations;
            stats8[1] = B.st.key_switches;
            stats8[2] = B.st.mults;
            stats8[3] = B.st.rescales;
            stats8[4] = B.st.plain_mults;
        }
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "btp_oracle_run: %s\n", e.what());
        return 1;
    }
}

// encoder probe: slots [n] (re, im interleaved) -> coefficient-domain limbs [level+1][N]
int btp_encode(int logN, const double* slots, int level, double scale, const u64* q, u64* out) {
    try {
        const int N = 1 << logN;
        lsb::btp::Encoder enc(N);
        std::vector<lsb::btp::cplx> v(N / 2);
        for (int i = 0; i < N / 2; i++) v[i] = lsb::btp::cplx(slots[2 * i], slots[2 * i + 1]);
        std::vector<u64> qq(q, q + level + 1);
        auto r = enc.encode(std::move(v), level, (long double)scale, qq);
        std::memcpy(out, r.data(), r.size() * sizeof(u64));
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "btp_encode: %s\n", e.what());
        return 1;
    }
}

void orc_gen_swk(void* h, const int8_t* s_in, const int8_t* s_out, u64 seed, int key_level, u64* evk) {
    const orc::Context& ctx = *(orc::Context*)h;
    orc::gen_switching_key(ctx, s_out, [&](int midx, u64* o) { orc::secret_ntt(ctx, midx, s_in, o); }, seed, key_level, evk);
}

}  // extern "C"
