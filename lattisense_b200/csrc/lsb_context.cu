// Context construction and cached table builders (host side).
#include "lsb_context.h"
#include "lsb_tables.h"
#include "lsb_ntt_cluster_core.h"
#include <cstring>

namespace lsb {

using namespace host;

Workspace::~Workspace() {
    if (buf_) {
        cudaSetDevice(device_);
        cudaFree(buf_);
    }
}
u64* Workspace::reserve(size_t words, cudaStream_t s) {
    if (words <= cap_) return buf_;
    LSB_CUDA(cudaSetDevice(device_));
    if (buf_) {
        LSB_CUDA(cudaStreamSynchronize(s));
        LSB_CUDA(cudaFree(buf_));
        buf_ = nullptr;
        cap_ = 0;
    }
    size_t want = words + words / 8;
    LSB_CUDA(cudaMalloc(&buf_, want * sizeof(u64)));
    cap_ = want;
    return buf_;
}

void* Context::to_device(const void* host_ptr, size_t bytes) {
    void* d = nullptr;
    LSB_CUDA(cudaSetDevice(device));
    LSB_CUDA(cudaMalloc(&d, bytes));
    LSB_CUDA(cudaMemcpy(d, host_ptr, bytes, cudaMemcpyHostToDevice));
    // a pageable H2D cudaMemcpy may return once the data sits in the driver's staging buffer, before the DMA
    // has landed; the kernels that read these tables run on non-blocking streams, which do not order
    // against the legacy stream, so wait for the DMA here (tables are built once and cached)
    LSB_CUDA(cudaStreamSynchronize(cudaStreamLegacy));
    owned_.push_back(d);
    return d;
}

Context::Context(int device_, int scheme_, int logN_, const u64* q, int nq_, const u64* p, int np_, u64 t_)
    : device(device_), scheme(scheme_), logN(logN_), N(1 << logN_), nq(nq_), np(np_), t(t_) {
    LSB_CHECK(logN >= 12 && logN <= 16, "lsb: ring degree must be 2^12..2^16");
    LSB_CHECK(nq >= 1 && np >= 1, "lsb: need at least one Q and one P modulus");
    LSB_CUDA(cudaSetDevice(device));
    for (int i = 0; i < nq; i++) moduli.push_back(q[i]);
    for (int j = 0; j < np; j++) moduli.push_back(p[j]);
    if (scheme == SCHEME_BFV) {
        // bfv.NewParameters / NewEvaluator (Lattigo v4): auxiliary basis QMul of 61-bit NTT primes,
        // levelQMul[i] = ceil((bitlen(Q_i) + logN) / 61) - 1.  Not carried by mega_ag.json -> regenerated.
        LSB_CHECK(t > 1, "lsb: BFV needs a plaintext modulus");
        BigU Q(1);
        level_qmul.resize(nq);
        for (int i = 0; i < nq; i++) {
            Q.mul(q[i]);
            level_qmul[i] = (Q.bits() + logN + 60) / 61 - 1;
        }
        nqm = (Q.bits() + logN + 60) / 61;
        for (u64 v : ntt_primes_below(61, 2ull << logN, nqm)) moduli.push_back(v);
    }
    const int nmod = (int)moduli.size();
    LSB_CHECK(nmod <= 250, "lsb: too many moduli");
    std::vector<Tw> fwd((size_t)nmod * N), inv((size_t)nmod * N);
    std::vector<TwD> fwd_d((size_t)nmod * N), inv_d((size_t)nmod * N);
    std::vector<ModConst> mc(nmod);
    for (int m = 0; m < nmod; m++) {
        build_ntt_tables(moduli[m], logN, &fwd[(size_t)m * N], &inv[(size_t)m * N], &mc[m], &fwd_d[(size_t)m * N],
                         &inv_d[(size_t)m * N]);
    }
    dt.device = device;
    dt.prof = &prof;
    cudaDeviceProp prop;
    LSB_CUDA(cudaGetDeviceProperties(&prop, device));
    dt.sm_count = prop.multiProcessorCount;
    dt.logN = logN;
    dt.N = N;
    dt.nmod = nmod;
    dt.tw_fwd = (const Tw*)to_device(fwd.data(), fwd.size() * sizeof(Tw));
    dt.tw_inv = (const Tw*)to_device(inv.data(), inv.size() * sizeof(Tw));
    dt.f64_ok.resize(nmod);
    for (int m = 0; m < nmod; m++) dt.f64_ok[m] = (uint8_t)mc[m].f64_ok;
    dt.elem_f64_ok.resize(nmod);
    for (int m = 0; m < nmod; m++) dt.elem_f64_ok[m] = moduli[m] < kElemF64Limit ? 1 : 0;
    dt.tw_fwd_d = (const TwD*)to_device(fwd_d.data(), fwd_d.size() * sizeof(TwD));
    dt.tw_inv_d = (const TwD*)to_device(inv_d.data(), inv_d.size() * sizeof(TwD));
    dt.mods = (const ModConst*)to_device(mc.data(), mc.size() * sizeof(ModConst));
    {
        const int lrl = cluster_last_lr(logN);
        for (int m = 0; m < nmod; m++) {
            soa_tail(&fwd[(size_t)m * N], logN, lrl);
            soa_tail(&inv[(size_t)m * N], logN, lrl);
            soa_tail(&fwd_d[(size_t)m * N], logN, lrl);
            soa_tail(&inv_d[(size_t)m * N], logN, lrl);
        }
        dt.twc_fwd = (const Tw*)to_device(fwd.data(), fwd.size() * sizeof(Tw));
        dt.twc_inv = (const Tw*)to_device(inv.data(), inv.size() * sizeof(Tw));
        dt.twc_fwd_d = (const TwD*)to_device(fwd_d.data(), fwd_d.size() * sizeof(TwD));
        dt.twc_inv_d = (const TwD*)to_device(inv_d.data(), inv_d.size() * sizeof(TwD));
        std::vector<double> w(fwd_d.size());
        for (size_t i = 0; i < w.size(); i++) w[i] = fwd_d[i].w;
        dt.twc_fwd_w = (const double*)to_device(w.data(), w.size() * sizeof(double));
        for (size_t i = 0; i < w.size(); i++) w[i] = inv_d[i].w;
        dt.twc_inv_w = (const double*)to_device(w.data(), w.size() * sizeof(double));
    }
    LSB_CUDA(cudaStreamCreateWithFlags(&dt.ntt_side, cudaStreamNonBlocking));
    LSB_CUDA(cudaEventCreateWithFlags(&dt.ntt_fork, cudaEventDisableTiming));
    LSB_CUDA(cudaEventCreateWithFlags(&dt.ntt_join, cudaEventDisableTiming));
    ws0_.reset(new Workspace(device));
}

Context::SidePipe& Context::side_pipe() {
    std::lock_guard<std::mutex> lk(mu_);
    if (!side_.stream) {
        LSB_CUDA(cudaSetDevice(device));
        LSB_CUDA(cudaStreamCreateWithFlags(&side_.stream, cudaStreamNonBlocking));
        LSB_CUDA(cudaEventCreateWithFlags(&side_.fork, cudaEventDisableTiming));
        LSB_CUDA(cudaEventCreateWithFlags(&side_.join, cudaEventDisableTiming));
    }
    return side_;
}

Context::~Context() {
    cudaSetDevice(device);
    if (side_.stream) {
        cudaStreamSynchronize(side_.stream);
        cudaStreamDestroy(side_.stream);
        cudaEventDestroy(side_.fork);
        cudaEventDestroy(side_.join);
    }
    if (dt.ntt_side) {
        cudaStreamSynchronize(dt.ntt_side);
        cudaStreamDestroy(dt.ntt_side);
        cudaEventDestroy(dt.ntt_fork);
        cudaEventDestroy(dt.ntt_join);
    }
    ws0_.reset();
    for (void* p : owned_) cudaFree(p);
}

// CRT tables for extending residues from the moduli `src` to each modulus of `dst`.
BexJob Context::make_bex(const std::vector<int>& src, const std::vector<int>& dst) const {
    BexJob j = make_bex_tables(moduli, src, dst);
    static const bool elem_f64 = [] {  // LSB_ELEM_F64=0: integer CRT everywhere (A/B switch)
        const char* e = getenv("LSB_ELEM_F64");
        return !(e && atoi(e) == 0);
    }();
    if (!elem_f64) j.f64_ok = 0;
    return j;
}

const void* Context::upload_jobs(const std::vector<BexJob>& jobs) {  // mu_ held
    const void* d = to_device(jobs.data(), jobs.size() * sizeof(BexJob));
    std::vector<BexMeta>& m = bex_meta_[d];
    for (const BexJob& j : jobs) m.push_back(BexMeta{j.ns, j.nt, j.mode, j.centered, j.f64_ok});
    return d;
}
const BexMeta* Context::bex_meta(const BexJob* dev) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = bex_meta_.find(dev);
    return it == bex_meta_.end() ? nullptr : it->second.data();
}

static inline u64 ckey(int kind, int a, int b = 0, int c = 0) {
    return ((u64)kind << 56) | ((u64)(a & 0xFFFFF) << 32) | ((u64)(b & 0xFFFF) << 16) | (u64)(c & 0xFFFF);
}

// Digits of a key-switch are extended and transformed in launch groups of at most this many: a launch addresses its rows
// through uint8_t slots (< 256) and at most kMaxRows of them.  Every parameter set up to beta * (l + k) = 192 -- all of
// BASELINE's, cfg5's 150 included -- is one group; the reference frontend's default CKKS n = 65536 set (30 Q + 4 P
// limbs, parameter.json) at its top levels (beta = 8, 272 extended limbs) takes two.
int Context::ks_digit_group(int level) const {
    const int l = level + 1, lk_ = l + np, b = beta(level);
    int g = b;
    while (g > 1 && (g * lk_ > 256 || g * lk_ - std::min(l, g * np) > kMaxRows)) g--;
    return g;
}

const BexJob* Context::modup_jobs(int level, int coeff_slot0) {
    std::lock_guard<std::mutex> lk(mu_);
    const u64 key = ckey(1, level, coeff_slot0);
    auto it = cache_.find(key);
    if (it != cache_.end()) return (const BexJob*)it->second;
    const int l = level + 1, alpha = np, b = beta(level), lk_ = l + np, grp = ks_digit_group(level);
    std::vector<BexJob> jobs;
    for (int d = 0; d < b; d++) {
        const int st = d * alpha, ed = std::min(st + alpha, l);
        const int dg = d % grp;  // destination slots are relative to the digit's launch group (ks_decompose, lsb_ops.cu)
        std::vector<int> src, dst, slots;
        for (int i = st; i < ed; i++) src.push_back(i);
        for (int i = 0; i < l; i++)
            if (i < st || i >= ed) {
                dst.push_back(i);
                slots.push_back(dg * lk_ + i);
            }
        for (int j = 0; j < np; j++) {
            dst.push_back(idxP(j));
            slots.push_back(dg * lk_ + l + j);
        }
        BexJob job = make_bex(src, dst);
        job.centered = (ed - st == 1) ? 1 : 0;  // rlwe DecomposeAndSplit single-modulus digit
        for (int i = 0; i < job.ns; i++) job.src_slot[i] = (uint8_t)(coeff_slot0 + st + i);
        LSB_CHECK(slots.back() < 256, "lsb: key-switch digit layout exceeds slot range");
        for (int tt = 0; tt < job.nt; tt++) job.dst_slot[tt] = (uint8_t)slots[tt];
        jobs.push_back(job);
    }
    const void* d = upload_jobs(jobs);
    cache_[key] = d;
    return (const BexJob*)d;
}

const BexJob* Context::moddown_jobs(int level) {
    std::lock_guard<std::mutex> lk(mu_);
    const u64 key = ckey(2, level);
    auto it = cache_.find(key);
    if (it != cache_.end()) return (const BexJob*)it->second;
    const int l = level + 1, lk_ = l + np;
    std::vector<int> src, dst;
    for (int j = 0; j < np; j++) src.push_back(idxP(j));
    for (int i = 0; i < l; i++) dst.push_back(i);
    std::vector<BexJob> jobs;
    for (int p = 0; p < 2; p++) {
        BexJob job = make_bex(src, dst);
        for (int j = 0; j < np; j++) job.src_slot[j] = (uint8_t)(p * lk_ + l + j);
        for (int i = 0; i < l; i++) job.dst_slot[i] = (uint8_t)(p * l + i);
        jobs.push_back(job);
    }
    const void* d = upload_jobs(jobs);
    cache_[key] = d;
    return (const BexJob*)d;
}

const FinishMap* Context::moddown_finish(int level, int addmask, int out_stride) {
    std::lock_guard<std::mutex> lk(mu_);
    const u64 key = ckey(3, level, addmask, out_stride);
    auto it = cache_.find(key);
    if (it != cache_.end()) return (const FinishMap*)it->second;
    const int l = level + 1, lk_ = l + np;
    FinishMap fm;
    std::memset(&fm, 0, sizeof(fm));
    fm.n = 2 * l;
    LSB_CHECK(fm.n <= kMaxRows, "lsb: too many limbs");
    for (int p = 0; p < 2; p++)
        for (int i = 0; i < l; i++) {
            const int r = p * l + i;
            const u64 qi = moduli[i];
            u64 pm = 1;
            for (int j = 0; j < np; j++) pm = mulmod(pm, moduli[idxP(j)] % qi, qi);
            fm.a_slot[r] = (uint8_t)(p * lk_ + i);
            fm.b_slot[r] = (uint8_t)(p * l + i);
            fm.add_slot[r] = ((addmask >> p) & 1) ? (uint8_t)(p * l + i) : (uint8_t)0xFF;
            fm.out_slot[r] = (uint8_t)(p * out_stride + i);
            fm.mod[r] = (uint8_t)i;
            fm.c[r] = make_tw(invmod(pm, qi), qi);
        }
    const void* d = to_device(&fm, sizeof(fm));
    cache_[key] = d;
    fm_host_[d] = fm;
    return (const FinishMap*)d;
}

const FinishMap* Context::rescale_finish(int level, int npoly) {
    std::lock_guard<std::mutex> lk(mu_);
    const u64 key = ckey(4, level, npoly);
    auto it = cache_.find(key);
    if (it != cache_.end()) return (const FinishMap*)it->second;
    const int l = level + 1;
    FinishMap fm;
    std::memset(&fm, 0, sizeof(fm));
    fm.n = npoly * level;
    LSB_CHECK(fm.n <= kMaxRows, "lsb: too many limbs");
    const u64 ql = moduli[level];
    for (int p = 0; p < npoly; p++)
        for (int i = 0; i < level; i++) {
            const int r = p * level + i;
            const u64 qi = moduli[i];
            fm.a_slot[r] = (uint8_t)(p * l + i);
            fm.b_slot[r] = (uint8_t)(p * level + i);
            fm.add_slot[r] = 0xFF;
            fm.out_slot[r] = (uint8_t)(p * (l - 1) + i);
            fm.mod[r] = (uint8_t)i;
            fm.c[r] = make_tw(invmod(ql % qi, qi), qi);
        }
    const void* d = to_device(&fm, sizeof(fm));
    cache_[key] = d;
    fm_host_[d] = fm;
    return (const FinishMap*)d;
}

const FinishMap* Context::finish_host(const FinishMap* dev) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = fm_host_.find(dev);
    LSB_CHECK(it != fm_host_.end(), "lsb: unknown finish map");
    return &it->second;
}

const Tw* Context::qlinv(int level) {
    std::lock_guard<std::mutex> lk(mu_);
    const u64 key = ckey(5, level);
    auto it = cache_.find(key);
    if (it != cache_.end()) return (const Tw*)it->second;
    std::vector<Tw> v(std::max(level, 1));
    const u64 ql = moduli[level];
    for (int i = 0; i < level; i++) v[i] = make_tw(invmod(ql % moduli[i], moduli[i]), moduli[i]);
    const void* d = to_device(v.data(), v.size() * sizeof(Tw));
    cache_[key] = d;
    return (const Tw*)d;
}

// ring.PermuteNTTIndex: out[i] = in[index[i]] realises X -> X^gal on a bit-reversed NTT vector
const u32* Context::permute_index(u64 gal) {
    std::lock_guard<std::mutex> lk(mu_);
    const u64 key = ckey(6, (int)(gal & 0xFFFFF), (int)(gal >> 20));
    auto it = cache_.find(key);
    if (it != cache_.end()) return (const u32*)it->second;
    std::vector<u32> idx(N);
    const u64 mask = 2ull * N - 1;
    for (u64 i = 0; i < (u64)N; i++) {
        const u64 t1 = 2 * bitrev(i, logN) + 1;
        const u64 t2 = ((gal * t1) & mask) - 1;
        idx[i] = (u32)bitrev(t2 >> 1, logN);
    }
    const void* d = to_device(idx.data(), idx.size() * sizeof(u32));
    cache_[key] = d;
    return (const u32*)d;
}

// ---- BFV multiply tables (bfv evaluator tensorAndRescale / quantize) ----
const BexJob* Context::bfv_q_to_qmul_jobs(int level) {
    std::lock_guard<std::mutex> lk(mu_);
    const u64 key = ckey(7, level);
    auto it = cache_.find(key);
    if (it != cache_.end()) return (const BexJob*)it->second;
    const int l = level + 1, lm = level_qmul[level] + 1, L = l + lm;
    std::vector<int> src, dst;
    for (int i = 0; i < l; i++) src.push_back(i);
    for (int j = 0; j < lm; j++) dst.push_back(idxQM(j));
    std::vector<BexJob> jobs;
    for (int p = 0; p < 2; p++) {
        BexJob job = make_bex(src, dst);
        for (int i = 0; i < l; i++) job.src_slot[i] = (uint8_t)(p * l + i);
        for (int j = 0; j < lm; j++) job.dst_slot[j] = (uint8_t)(p * L + l + j);
        jobs.push_back(job);
    }
    const void* d = upload_jobs(jobs);
    cache_[key] = d;
    return (const BexJob*)d;
}

const BexJob* Context::bfv_quant1_jobs(int level) {
    std::lock_guard<std::mutex> lk(mu_);
    const u64 key = ckey(8, level);
    auto it = cache_.find(key);
    if (it != cache_.end()) return (const BexJob*)it->second;
    const int l = level + 1, lm = level_qmul[level] + 1, L = l + lm;
    BigU QM(1);
    for (int j = 0; j < lm; j++) QM.mul(moduli[idxQM(j)]);
    BigU half = QM;
    half.shr1();
    std::vector<int> src, dst;
    for (int i = 0; i < l; i++) src.push_back(i);
    for (int j = 0; j < lm; j++) dst.push_back(idxQM(j));
    std::vector<BexJob> jobs;
    for (int p = 0; p < 3; p++) {
        // ModDownQPtoP: (x_QMul - ext([x]_Q)) * Q^-1, then + QMul/2
        BexJob job = make_bex(src, dst);
        job.mode = BEX_OTHER_SUB;
        for (int i = 0; i < l; i++) job.src_slot[i] = (uint8_t)(p * L + i);
        for (int j = 0; j < lm; j++) {
            const u64 m = moduli[idxQM(j)];
            u64 qprod = 1;
            for (int i = 0; i < l; i++) qprod = mulmod(qprod, moduli[i] % m, m);
            job.dst_slot[j] = (uint8_t)(p * lm + j);
            job.other_slot[j] = (uint8_t)(p * L + l + j);
            job.mulc[j] = make_tw(invmod(qprod, m), m);
            job.addc[j] = half.mod(m);
        }
        bex_fill_f64(job, moduli);
        jobs.push_back(job);
    }
    const void* d = upload_jobs(jobs);
    cache_[key] = d;
    return (const BexJob*)d;
}

const BexJob* Context::bfv_quant2_jobs(int level) {
    std::lock_guard<std::mutex> lk(mu_);
    const u64 key = ckey(9, level);
    auto it = cache_.find(key);
    if (it != cache_.end()) return (const BexJob*)it->second;
    const int l = level + 1, lm = level_qmul[level] + 1;
    BigU QM(1);
    for (int j = 0; j < lm; j++) QM.mul(moduli[idxQM(j)]);
    BigU half = QM;
    half.shr1();
    std::vector<int> src, dst;
    for (int j = 0; j < lm; j++) src.push_back(idxQM(j));
    for (int i = 0; i < l; i++) dst.push_back(i);
    std::vector<BexJob> jobs;
    for (int p = 0; p < 3; p++) {
        // ModUpPtoQ, subtract QMul/2, multiply by t
        BexJob job = make_bex(src, dst);
        job.mode = BEX_SUB_MUL;
        for (int j = 0; j < lm; j++) job.src_slot[j] = (uint8_t)(p * lm + j);
        for (int i = 0; i < l; i++) {
            const u64 qi = moduli[i];
            job.dst_slot[i] = (uint8_t)(p * l + i);
            job.addc[i] = half.mod(qi);
            job.mulc[i] = make_tw(t % qi, qi);
        }
        bex_fill_f64(job, moduli);
        jobs.push_back(job);
    }
    const void* d = upload_jobs(jobs);
    cache_[key] = d;
    return (const BexJob*)d;
}

// bfv Scaler.ScaleUp (T coprime with Q): pt_i = -(Q*m + t/2 mod t, centred) * t^-1 mod q_i
const ScaleUpConsts* Context::bfv_scale_up_consts(int level) {
    std::lock_guard<std::mutex> lk(mu_);
    const u64 key = ckey(10, level);
    auto it = cache_.find(key);
    if (it != cache_.end()) return (const ScaleUpConsts*)it->second;
    const int l = level + 1;
    LSB_CHECK(l <= 64, "lsb: too many limbs");
    ScaleUpConsts sc;
    std::memset(&sc, 0, sizeof(sc));
    sc.t = t;
    sc.thalf = t >> 1;
    sc.bt = make_barrett(t);
    u64 qmodt = 1;
    for (int i = 0; i < l; i++) qmodt = mulmod(qmodt, moduli[i] % t, t);
    sc.qmodt = qmodt;
    for (int i = 0; i < l; i++) {
        const u64 qi = moduli[i];
        const u64 tinv = invmod(t % qi, qi);
        sc.thalf_q[i] = sc.thalf % qi;
        sc.neg_tinv[i] = make_tw(tinv ? qi - tinv : 0, qi);
    }
    const void* d = to_device(&sc, sizeof(sc));
    cache_[key] = d;
    return (const ScaleUpConsts*)d;
}

}  // namespace lsb
