// CKKS bootstrapping on the device: the pipeline of lsb_bootstrap.h instantiated with the batched device operations of
// lsb_ops.h.  Replaces bind_gpu_bootstrap (mega_ag_runners/gpu/mega_ag_executors_gpu.cu:410-426 ->
// HEonGPU regular_bootstrapping_v2) and the per-run bootstrapping precompute of init_gpu_context
// (mega_ag_runners/gpu/gpu_wrapper.cu:106-116): here the CoeffsToSlots / SlotsToCoeffs plaintexts are encoded once per
// (task, device) and stay resident in HBM.
#include <cstring>
#include <mutex>

#include "lsb_bootstrap_dev.h"
#include "lsb_host_math.h"

namespace lsb {

namespace {

struct DevOps {
    struct Ct {
        u64* p = nullptr;
        int level = -1;
        int nb = 0;  // batch items held (EvalMod runs on the real and imaginary halves stacked into one batch)
        cudaStream_t s = nullptr;
        Ct() = default;
        Ct(u64* p_, int level_, int nb_, cudaStream_t s_) : p(p_), level(level_), nb(nb_), s(s_) {}
        Ct(const Ct&) = delete;
        Ct& operator=(const Ct&) = delete;
        Ct(Ct&& o) noexcept : p(o.p), level(o.level), nb(o.nb), s(o.s) { o.p = nullptr; }
        Ct& operator=(Ct&& o) noexcept {
            if (this != &o) {
                release();
                p = o.p;
                level = o.level;
                nb = o.nb;
                s = o.s;
                o.p = nullptr;
            }
            return *this;
        }
        ~Ct() { release(); }
        void release() {
            if (p) cudaFreeAsync(p, s);  // stream-ordered: every reader was enqueued on s before the handle died
            p = nullptr;
        }
    };
    struct Pt {
        const u64* p = nullptr;  // [level+1][N] NTT domain; owned by the DeviceBootstrap
        int level = -1;
    };

    OpEnv e;
    int batch;
    const BootstrapKeys& keys;
    std::vector<void*>& owned;  // persistent allocations (plaintexts) of the DeviceBootstrap
    const size_t N;
    long long* launches_hint = nullptr;

    DevOps(OpEnv env, int batch_, const BootstrapKeys& k, std::vector<void*>& own) : e(env), batch(batch_), keys(k), owned(own), N(env.c.N) {}

    int level(const Ct& a) const { return a.level; }
    size_t words(int level) const { return (size_t)2 * (level + 1) * N; }
    Ct make(int level, int nb) {
        u64* p = nullptr;
        LSB_CUDA(cudaMallocAsync((void**)&p, words(level) * nb * sizeof(u64), e.s));
        return Ct(p, level, nb, e.s);
    }
    Ct concat(const Ct& a, const Ct& b) {
        LSB_CHECK(a.level == b.level, "lsb bootstrap: concat of different levels");
        Ct o = make(a.level, a.nb + b.nb);
        const size_t wa = words(a.level) * a.nb, wb = words(a.level) * b.nb;
        LSB_CUDA(cudaMemcpyAsync(o.p, a.p, wa * sizeof(u64), cudaMemcpyDeviceToDevice, e.s));
        LSB_CUDA(cudaMemcpyAsync(o.p + wa, b.p, wb * sizeof(u64), cudaMemcpyDeviceToDevice, e.s));
        return o;
    }
    std::pair<Ct, Ct> split(const Ct& c) {
        Ct a = make(c.level, c.nb / 2), b = make(c.level, c.nb - c.nb / 2);
        const size_t wa = words(c.level) * a.nb, wb = words(c.level) * b.nb;
        LSB_CUDA(cudaMemcpyAsync(a.p, c.p, wa * sizeof(u64), cudaMemcpyDeviceToDevice, e.s));
        LSB_CUDA(cudaMemcpyAsync(b.p, c.p + wa, wb * sizeof(u64), cudaMemcpyDeviceToDevice, e.s));
        return {std::move(a), std::move(b)};
    }
    BatchPtr bp(const Ct& a) const { return bp_strided(a.p, words(a.level)); }

    Pt make_plain(int level, const std::vector<u64>& coeff) {
        LSB_CHECK(coeff.size() == (size_t)(level + 1) * N, "lsb bootstrap: plaintext size mismatch");
        u64* d = nullptr;
        LSB_CUDA(cudaSetDevice(e.c.device));
        LSB_CUDA(cudaMalloc((void**)&d, coeff.size() * sizeof(u64)));
        owned.push_back(d);
        LSB_CUDA(cudaMemcpyAsync(d, coeff.data(), coeff.size() * sizeof(u64), cudaMemcpyHostToDevice, e.s));
        std::vector<int> mi(level + 1);
        for (int i = 0; i <= level; i++) mi[i] = i;
        op_ntt(e, bp_strided(d, coeff.size()), bp_strided(d, coeff.size()), level + 1, mi.data(), level + 1, false, 1);
        // the host vector dies when the caller returns: pageable H2D copies are staged before cudaMemcpyAsync returns
        Pt p;
        p.p = d;
        p.level = level;
        return p;
    }
    u64 gal_of(int step) const {
        const u64 m = 2 * (u64)N;
        u64 r = 1, b = 5 % m;
        for (u64 x = (u64)step; x; x >>= 1, b = b * b % m)
            if (x & 1) r = r * b % m;
        return r;
    }
    Ct rotate_gal(const Ct& a, u64 gal) {
        const u64* key = nullptr;
        for (size_t i = 0; i < keys.gal.size(); i++)
            if (keys.gal[i] == gal) key = keys.glk[i];
        if (!key) throw std::runtime_error("lsb bootstrap: the task carries no Galois key for element " + std::to_string(gal));
        Ct o = make(a.level, a.nb);
        op_rotate(e, a.level, bp(a), key, keys.key_level, gal, bp(o), a.nb);
        return o;
    }
    Ct rotate(const Ct& a, int step) { return rotate_gal(a, gal_of(step)); }
    const u64* key_of(u64 gal) const {
        for (size_t i = 0; i < keys.gal.size(); i++)
            if (keys.gal[i] == gal) return keys.glk[i];
        throw std::runtime_error("lsb bootstrap: the task carries no Galois key for element " + std::to_string(gal));
    }
    std::vector<Ct> rotate_many(const Ct& a, const std::vector<int>& steps) {
        static const bool hoist = [] {  // LSB_BTP_HOIST=0: one full key-switch per rotation (A/B; limbs are identical)
            const char* e = getenv("LSB_BTP_HOIST");
            return !(e && atoi(e) == 0);
        }();
        std::vector<Ct> out;
        if (!hoist || steps.size() < 2) {
            for (int st : steps) out.push_back(rotate(a, st));
            return out;
        }
        std::vector<const u64*> ks;
        std::vector<u64> gals;
        std::vector<BatchPtr> outs;
        for (int st : steps) {
            gals.push_back(gal_of(st));
            ks.push_back(key_of(gals.back()));
            out.push_back(make(a.level, a.nb));
            outs.push_back(bp(out.back()));
        }
        op_rotate_hoisted(e, a.level, bp(a), (int)steps.size(), ks.data(), keys.key_level, gals.data(), outs.data(), a.nb);
        return out;
    }
    Ct conjugate(const Ct& a) { return rotate_gal(a, 2 * (u64)N - 1); }
    Ct switch_key(const Ct& a, int which) {
        Ct o = make(a.level, a.nb);
        op_switch_key(e, a.level, bp(a), which == 0 ? keys.swk_dts : keys.swk_std, which == 0 ? keys.swk_dts_level : keys.key_level, bp(o),
                      a.nb);
        return o;
    }
    Ct mod_raise(const Ct& a, int top) {
        LSB_CHECK(a.level == 0, "lsb bootstrap: mod_raise expects a level-0 ciphertext");
        Ct o = make(top, a.nb);
        op_mod_raise(e, top, 2, bp(a), bp(o), a.nb);
        return o;
    }
    Ct add(const Ct& a, const Ct& b) {
        Ct o = make(a.level, a.nb);
        op_add(e, a.level, 2, bp(a), bp(b), bp(o), a.nb);
        return o;
    }
    Ct sub(const Ct& a, const Ct& b) {
        Ct o = make(a.level, a.nb);
        op_sub(e, a.level, 2, bp(a), bp(b), bp(o), a.nb);
        return o;
    }
    Ct mul_relin(const Ct& a, const Ct& b) {
        LSB_CHECK(a.level == b.level, "lsb bootstrap: ct x ct operands at different levels");
        Ct o = make(a.level, a.nb);
        op_mult_relin(e, a.level, bp(a), bp(b), keys.rlk, keys.key_level, bp(o), a.nb);
        return o;
    }
    Ct rescale(const Ct& a) {
        Ct o = make(a.level - 1, a.nb);
        op_rescale(e, a.level, 2, bp(a), bp(o), a.nb);
        return o;
    }
    Ct drop_to(const Ct& a, int level) {
        Ct o = make(level, a.nb);
        op_drop_to(e, a.level, level, 2, bp(a), bp(o), a.nb);
        return o;
    }
    Ct mul_scalar(const Ct& a, const std::vector<u64>& k) {
        Ct o = make(a.level, a.nb);
        op_scalar(e, a.level, 2, bp(a), k.data(), false, bp(o), a.nb);
        return o;
    }
    Ct add_scalar(const Ct& a, const std::vector<u64>& k) {
        Ct o = make(a.level, a.nb);
        op_scalar(e, a.level, 2, bp(a), k.data(), true, bp(o), a.nb);
        return o;
    }
    // pointer table [term][batch] on the device; the table itself is freed in stream order
    const u64** upload_table(const std::vector<const u64*>& v) {
        const u64** d = nullptr;
        LSB_CUDA(cudaMallocAsync((void**)&d, v.size() * sizeof(u64*), e.s));
        LSB_CUDA(cudaMemcpyAsync(d, v.data(), v.size() * sizeof(u64*), cudaMemcpyHostToDevice, e.s));
        return d;
    }
    Ct sum_terms(const std::vector<const Ct*>& cts, const std::vector<const Pt*>& pts) {
        const int lv = cts[0]->level, n_mac = (int)pts.size(), n_all = (int)cts.size(), nbat = cts[0]->nb;
        Ct o = make(lv, nbat);
        std::vector<const u64*> tab((size_t)(n_all + n_mac + 1) * nbat);
        for (int t = 0; t < n_all; t++) {
            LSB_CHECK(cts[t]->level == lv, "lsb bootstrap: sum operands at different levels");
            for (int b = 0; b < nbat; b++) tab[(size_t)t * nbat + b] = cts[t]->p + (size_t)b * words(lv);
        }
        for (int t = 0; t < n_mac; t++) {
            LSB_CHECK(pts[t]->level == lv, "lsb bootstrap: plaintext at the wrong level");
            for (int b = 0; b < nbat; b++) tab[(size_t)(n_all + t) * nbat + b] = pts[t]->p;
        }
        for (int b = 0; b < nbat; b++) tab[(size_t)(n_all + n_mac) * nbat + b] = o.p + (size_t)b * words(lv);
        const u64** d = upload_table(tab);
        op_sum_terms(e, lv, 2, d, d + (size_t)n_all * nbat, (u64* const*)(d + (size_t)(n_all + n_mac) * nbat), n_mac, n_all - n_mac, nbat);
        LSB_CUDA(cudaFreeAsync((void*)d, e.s));
        return o;
    }
    Ct sum_products(const std::vector<const Ct*>& cts, const std::vector<const Pt*>& pts) { return sum_terms(cts, pts); }
    Ct mul_plain(const Ct& a, const Pt& p) { return sum_terms({&a}, {&p}); }
    Ct sum(const std::vector<const Ct*>& cts) { return sum_terms(cts, {}); }
};

}  // namespace

struct DeviceBootstrap::Impl {
    btp::Params P;
    btp::Bootstrapper<DevOps> B;
    std::vector<void*> owned;
    std::mutex mu;
    explicit Impl(const btp::Params& p) : P(p), B(p) {}
};

DeviceBootstrap::DeviceBootstrap(Context& c, const btp::Params& p) : ctx_(c), impl_(new Impl(p)) {
    LSB_CHECK(c.scheme == SCHEME_CKKS, "lsb bootstrap: CKKS only");
    LSB_CHECK(p.logN == c.logN && (int)p.q.size() == c.nq && (int)p.p.size() == c.np, "lsb bootstrap: parameters do not match the context");
}
DeviceBootstrap::~DeviceBootstrap() {
    cudaSetDevice(ctx_.device);
    for (void* p : impl_->owned) cudaFree(p);
}
std::vector<int> DeviceBootstrap::rotations() const { return impl_->B.rotations(); }
int DeviceBootstrap::out_level() const { return impl_->P.out_level; }

void DeviceBootstrap::run(Workspace& ws, cudaStream_t s, const BootstrapKeys& keys, BatchPtr in, BatchPtr out, int batch, BootstrapStats* st) {
    std::lock_guard<std::mutex> lk(impl_->mu);
    LSB_CUDA(cudaSetDevice(ctx_.device));
    OpEnv e{ctx_, ws, s, 0};
    DevOps ops(e, batch, keys, impl_->owned);
    // gather the batch (independently allocated level-0 ciphertexts) into one block
    DevOps::Ct ct0 = ops.make(0, batch);
    op_drop_to(e, 0, 0, 2, in, ops.bp(ct0), batch);
    impl_->B.st = {};
    DevOps::Ct res = impl_->B.run(ops, ct0);
    LSB_CHECK(res.level == impl_->P.out_level, "lsb bootstrap: unexpected output level");
    op_drop_to(e, res.level, res.level, 2, ops.bp(res), out, batch);
    if (st) {
        st->rotations = impl_->B.st.rotations;
        st->key_switches = impl_->B.st.key_switches + impl_->B.st.rotations;
        st->mults = impl_->B.st.mults;
        st->rescales = impl_->B.st.rescales;
        st->plain_mults = impl_->B.st.plain_mults;
    }
}

}  // namespace lsb
