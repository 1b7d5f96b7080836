// extern "C" surface declared in include/lsb200.h.
#include <mutex>
#include <cstring>
#include "../../include/lsb200.h"
#include "lsb_ops.h"
#include "lsb_bootstrap_dev.h"

using namespace lsb;

struct lsb_context_st {
    std::unique_ptr<Context> c;
    std::mutex mu;  // serialises users of the default workspace
    int chunk = 0;
    // the default workspace is shared by every call on this context: when a call arrives on another stream than the
    // previous one, it is ordered after the previous stream's work (which may still be using the scratch block)
    cudaStream_t last_stream = nullptr;
    bool has_last = false;
    cudaEvent_t switch_ev = nullptr;
    ~lsb_context_st() {
        if (switch_ev && c) {
            cudaSetDevice(c->device);
            cudaEventDestroy(switch_ev);
        }
    }
};

static void order_after_previous_stream(lsb_context_st* h, cudaStream_t s) {  // h->mu held
    if (h->has_last && h->last_stream != s) {
        LSB_CUDA(cudaSetDevice(h->c->device));
        if (!h->switch_ev) LSB_CUDA(cudaEventCreateWithFlags(&h->switch_ev, cudaEventDisableTiming));
        if (cudaEventRecord(h->switch_ev, h->last_stream) == cudaSuccess) {
            LSB_CUDA(cudaStreamWaitEvent(s, h->switch_ev, 0));
        } else {  // the previous stream no longer exists: everything it was given has to be complete before the scratch is reused
            cudaGetLastError();
            LSB_CUDA(cudaDeviceSynchronize());
        }
    }
    h->last_stream = s;
    h->has_last = true;
}

static thread_local std::string g_err;

Context& lsb_context_ref(lsb_context_t h) {  // for the other translation units behind lsb200.h (lsb_shard_nccl.cu)
    LSB_CHECK(h && h->c, "lsb: null context");
    return *h->c;
}

void lsb_set_last_error(const std::string& what) { g_err = what; }  // for lsb_noalloc.cu

#define LSB_API_BEGIN \
    g_err.clear();    \
    try {
#define LSB_API_END(fail)                 \
    }                                     \
    catch (const std::exception& ex) {    \
        g_err = ex.what();                \
        return fail;                      \
    }                                     \
    catch (...) {                         \
        g_err = "lsb: unknown exception"; \
        return fail;                      \
    }

template <class F>
static int run_op(lsb_context_t h, void* stream, F&& f) {
    LSB_API_BEGIN
    LSB_CHECK(h && h->c, "lsb: null context");
    std::lock_guard<std::mutex> lk(h->mu);
    order_after_previous_stream(h, (cudaStream_t)stream);
    OpEnv e{*h->c, h->c->default_workspace(), (cudaStream_t)stream, h->chunk};
    f(e);
    return 0;
    LSB_API_END(1)
}

static BatchPtr dense(const uint64_t* p, size_t words) { return bp_strided(p, words); }

extern "C" {

const char* lsb_last_error(void) { return g_err.c_str(); }
const char* lsb_version(void) { return "lattisense-b200 0.1 (sm_100a)"; }

lsb_context_t lsb_context_create(int device, int scheme, int log_n, const uint64_t* q, int nq, const uint64_t* p,
                                 int np, uint64_t t) {
    LSB_API_BEGIN
    int ndev = 0;
    cudaError_t err = cudaGetDeviceCount(&ndev);
    if (err != cudaSuccess || ndev == 0)
        throw std::runtime_error(std::string("lsb: no CUDA device available (") + cudaGetErrorString(err) +
                                 "); this library has no CPU fallback");
    LSB_CHECK(device >= 0 && device < ndev, "lsb: device index out of range");
    auto* h = new lsb_context_st();
    try {
        h->c.reset(new Context(device, scheme, log_n, q, nq, p, np, t));
    } catch (...) {
        delete h;
        throw;
    }
    return h;
    LSB_API_END(nullptr)
}
void lsb_context_destroy(lsb_context_t h) { delete h; }
int lsb_context_num_moduli(lsb_context_t h) { return h && h->c ? (int)h->c->moduli.size() : 0; }
uint64_t lsb_context_modulus(lsb_context_t h, int i) {
    return (h && h->c && i >= 0 && i < (int)h->c->moduli.size()) ? h->c->moduli[i] : 0;
}
int lsb_context_level_qmul(lsb_context_t h, int level) {
    return (h && h->c && level >= 0 && level < (int)h->c->level_qmul.size()) ? h->c->level_qmul[level] : -1;
}
void lsb_set_chunk(lsb_context_t h, int chunk) {
    if (h) h->chunk = chunk;
}

uint64_t* lsb_device_alloc(lsb_context_t h, size_t words) {
    LSB_API_BEGIN
    LSB_CHECK(h && h->c, "lsb: null context");
    LSB_CUDA(cudaSetDevice(h->c->device));
    void* p = nullptr;
    LSB_CUDA(cudaMalloc(&p, words * sizeof(uint64_t)));
    return (uint64_t*)p;
    LSB_API_END(nullptr)
}
void lsb_device_free(lsb_context_t h, uint64_t* d) {
    if (!h || !h->c || !d) return;
    cudaSetDevice(h->c->device);
    cudaFree(d);
}
int lsb_upload(lsb_context_t h, uint64_t* d_dst, const uint64_t* h_src, size_t words, void* stream) {
    LSB_API_BEGIN
    LSB_CHECK(h && h->c, "lsb: null context");
    LSB_CUDA(cudaSetDevice(h->c->device));
    LSB_CUDA(cudaMemcpyAsync(d_dst, h_src, words * 8, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return 0;
    LSB_API_END(1)
}
int lsb_download(lsb_context_t h, uint64_t* h_dst, const uint64_t* d_src, size_t words, void* stream) {
    LSB_API_BEGIN
    LSB_CHECK(h && h->c, "lsb: null context");
    LSB_CUDA(cudaSetDevice(h->c->device));
    LSB_CUDA(cudaMemcpyAsync(h_dst, d_src, words * 8, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return 0;
    LSB_API_END(1)
}
int lsb_stream_sync(lsb_context_t h, void* stream) {
    LSB_API_BEGIN
    LSB_CHECK(h && h->c, "lsb: null context");
    LSB_CUDA(cudaSetDevice(h->c->device));
    LSB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
    LSB_API_END(1)
}

int lsb_ntt(lsb_context_t h, const uint64_t* d_in, uint64_t* d_out, int rows, const int* mod_idx, int period,
            int inverse, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const size_t w = (size_t)rows * e.c.N;
        op_ntt(e, dense(d_in, w), dense(d_out, w), rows, mod_idx, period, inverse != 0, batch);
    });
}
int lsb_mult(lsb_context_t h, int level, const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out3, int batch,
             void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        op_mult(e, level, dense(d_a, e.c.ct_words(level, 2)), dense(d_b, e.c.ct_words(level, 2)),
                dense(d_out3, e.c.ct_words(level, 3)), batch);
    });
}
int lsb_relin(lsb_context_t h, int level, const uint64_t* d_ct3, const uint64_t* d_rlk, int key_level,
              uint64_t* d_out2, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        op_relin(e, level, dense(d_ct3, e.c.ct_words(level, 3)), d_rlk, key_level,
                 dense(d_out2, e.c.ct_words(level, 2)), batch);
    });
}
int lsb_rescale(lsb_context_t h, int level, int npoly, const uint64_t* d_in, uint64_t* d_out, int batch,
                void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        op_rescale(e, level, npoly, dense(d_in, e.c.ct_words(level, npoly)),
                   dense(d_out, e.c.ct_words(level - 1, npoly)), batch);
    });
}
int lsb_mod_raise(lsb_context_t h, int top_level, int npoly, const uint64_t* d_in, uint64_t* d_out, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        op_mod_raise(e, top_level, npoly, dense(d_in, e.c.ct_words(0, npoly)), dense(d_out, e.c.ct_words(top_level, npoly)), batch);
    });
}
int lsb_linear_transform(lsb_context_t h, int level, const uint64_t* d_ct, int n_diag, const int* diag_rot,
                         const uint64_t* const* d_diag_pt, int n1, int n_keys, const int* key_rot, const uint64_t* const* d_glk,
                         int key_level, uint64_t* d_out, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        op_linear_transform(e, level, d_ct, n_diag, diag_rot, d_diag_pt, n1, n_keys, key_rot, d_glk, key_level, d_out);
    });
}
int lsb_drop_level(lsb_context_t h, int level, int npoly, const uint64_t* d_in, uint64_t* d_out, int batch,
                   void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        op_drop_level(e, level, npoly, dense(d_in, e.c.ct_words(level, npoly)),
                      dense(d_out, e.c.ct_words(level - 1, npoly)), batch);
    });
}
int lsb_rotate(lsb_context_t h, int level, const uint64_t* d_ct, const uint64_t* d_glk, int key_level,
               uint64_t galois_element, uint64_t* d_out, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        op_rotate(e, level, dense(d_ct, e.c.ct_words(level, 2)), d_glk, key_level, galois_element,
                  dense(d_out, e.c.ct_words(level, 2)), batch);
    });
}
int lsb_add(lsb_context_t h, int level, int npoly, const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out,
            int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const size_t w = e.c.ct_words(level, npoly);
        op_add(e, level, npoly, dense(d_a, w), dense(d_b, w), dense(d_out, w), batch);
    });
}
int lsb_sub(lsb_context_t h, int level, int npoly, const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out,
            int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const size_t w = e.c.ct_words(level, npoly);
        op_sub(e, level, npoly, dense(d_a, w), dense(d_b, w), dense(d_out, w), batch);
    });
}
int lsb_neg(lsb_context_t h, int level, int npoly, const uint64_t* d_a, uint64_t* d_out, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const size_t w = e.c.ct_words(level, npoly);
        op_neg(e, level, npoly, dense(d_a, w), dense(d_out, w), batch);
    });
}
int lsb_add_plain(lsb_context_t h, int level, int npoly, const uint64_t* d_ct, const uint64_t* d_pt,
                  uint64_t* d_out, int subtract, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const size_t w = e.c.ct_words(level, npoly);
        op_add_plain(e, level, npoly, dense(d_ct, w), dense(d_pt, e.c.ct_words(level, 1)), dense(d_out, w),
                     subtract != 0, batch);
    });
}
int lsb_add_plain_ringt(lsb_context_t h, int level, int npoly, const uint64_t* d_ct, const uint64_t* d_pt,
                        uint64_t* d_out, int subtract, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const size_t w = e.c.ct_words(level, npoly);
        op_add_plain_ringt(e, level, npoly, dense(d_ct, w), dense(d_pt, e.c.N), dense(d_out, w), subtract != 0,
                           batch);
    });
}
int lsb_mul_plain(lsb_context_t h, int level, int npoly, const uint64_t* d_ct, const uint64_t* d_pt,
                  uint64_t* d_out, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const size_t w = e.c.ct_words(level, npoly);
        op_mul_plain(e, level, npoly, dense(d_ct, w), dense(d_pt, e.c.ct_words(level, 1)), dense(d_out, w), batch);
    });
}
int lsb_mul_plain_ringt(lsb_context_t h, int level, int npoly, const uint64_t* d_ct, const uint64_t* d_pt,
                        uint64_t* d_out, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const size_t w = e.c.ct_words(level, npoly);
        op_mul_plain_ringt(e, level, npoly, dense(d_ct, w), dense(d_pt, e.c.N), dense(d_out, w), batch);
    });
}
int lsb_mult_relin(lsb_context_t h, int level, const uint64_t* d_a, const uint64_t* d_b, const uint64_t* d_rlk,
                   int key_level, uint64_t* d_out2, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const size_t w = e.c.ct_words(level, 2);
        op_mult_relin(e, level, dense(d_a, w), dense(d_b, w), d_rlk, key_level, dense(d_out2, w), batch);
    });
}
int lsb_mult_relin_rescale(lsb_context_t h, int level, const uint64_t* d_a, const uint64_t* d_b,
                           const uint64_t* d_rlk, int key_level, uint64_t* d_out2, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const size_t w = e.c.ct_words(level, 2);
        op_mult_relin_rescale(e, level, dense(d_a, w), dense(d_b, w), d_rlk, key_level,
                              dense(d_out2, e.c.ct_words(level - 1, 2)), batch);
    });
}

long long lsb_launch_count(lsb_context_t h) { return h && h->c ? h->c->prof.launches.load() : 0; }
void lsb_profile_enable(lsb_context_t h, int on) {
    if (!h || !h->c) return;
    h->c->prof.reset();
    h->c->prof.enabled.store(on != 0);
}
int lsb_profile_read(lsb_context_t h, int kind, double* ms, long long* kernels, double* algorithmic_bytes) {
    LSB_API_BEGIN
    LSB_CHECK(h && h->c && kind >= 0 && kind < PK_COUNT, "lsb: bad profile query");
    LSB_CUDA(cudaSetDevice(h->c->device));
    h->c->prof.read(kind, ms, kernels, algorithmic_bytes);
    return 0;
    LSB_API_END(1)
}


// ---- limb-sharded key-switch (include/lsb200.h, "multi-GPU" section) ----
int lsb_ks_shard_layout(lsb_context_t h, int level, int world, int rank, int* nq_own, int* np_own, int* cq, int* cp,
                        int* own_q, int* own_p) {
    LSB_API_BEGIN
    LSB_CHECK(h && h->c, "lsb: null context");
    const ShardPlan& sp = h->c->shard_plan(level, world, rank);
    if (nq_own) *nq_own = sp.nq_own;
    if (np_own) *np_own = sp.np_own;
    if (cq) *cq = sp.cq;
    if (cp) *cp = sp.cp;
    if (own_q)
        for (int u = 0; u < sp.nq_own; u++) own_q[u] = sp.own_q[u];
    if (own_p)
        for (int u = 0; u < sp.np_own; u++) own_p[u] = sp.own_p[u];
    return 0;
    LSB_API_END(1)
}
int lsb_ks_shard_begin(lsb_context_t h, int level, int world, int rank, const uint64_t* d_c_own_ntt, uint64_t* d_coeff_own,
                       int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const ShardPlan& sp = e.c.shard_plan(level, world, rank);
        op_ks_shard_begin(e, sp, dense(d_c_own_ntt, (size_t)sp.nq_own * e.c.N), dense(d_coeff_own, (size_t)sp.cq * e.c.N), batch);
    });
}
int lsb_ks_shard_inner(lsb_context_t h, int level, int world, int rank, const uint64_t* d_coeff_all,
                       const uint64_t* d_c_own_ntt, const uint64_t* d_evk_own, uint64_t* d_acc_own, uint64_t* d_accP_own,
                       int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const ShardPlan& sp = e.c.shard_plan(level, world, rank);
        const size_t N = e.c.N;
        op_ks_shard_inner(e, sp, dense(d_coeff_all, (size_t)world * sp.cq * N), dense(d_c_own_ntt, (size_t)sp.nq_own * N),
                          d_evk_own, dense(d_acc_own, (size_t)2 * sp.n_own * N), dense(d_accP_own, (size_t)2 * sp.cp * N), batch);
    });
}
int lsb_ks_shard_begin_p2p(lsb_context_t h, int level, int world, int rank, const uint64_t* d_c_own_ntt,
                           uint64_t* const* peer_coeff_all, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const ShardPlan& sp = e.c.shard_plan(level, world, rank);
        LSB_CHECK(peer_coeff_all, "lsb: null peer table");
        op_ks_shard_begin(e, sp, dense(d_c_own_ntt, (size_t)sp.nq_own * e.c.N), dense(d_c_own_ntt, 0), 1, peer_coeff_all);
    });
}
int lsb_ks_shard_inner_p2p(lsb_context_t h, int level, int world, int rank, const uint64_t* d_coeff_all,
                           const uint64_t* d_c_own_ntt, const uint64_t* d_evk_own, uint64_t* d_acc_own,
                           uint64_t* const* peer_accP_all, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const ShardPlan& sp = e.c.shard_plan(level, world, rank);
        const size_t N = e.c.N;
        LSB_CHECK(peer_accP_all, "lsb: null peer table");
        op_ks_shard_inner(e, sp, dense(d_coeff_all, (size_t)world * sp.cq * N), dense(d_c_own_ntt, (size_t)sp.nq_own * N),
                          d_evk_own, dense(d_acc_own, (size_t)2 * sp.n_own * N), dense(d_acc_own, 0), 1, peer_accP_all);
    });
}
int lsb_ks_shard_end(lsb_context_t h, int level, int world, int rank, const uint64_t* d_accP_all, const uint64_t* d_acc_own,
                     const uint64_t* d_addend_own, uint64_t* d_out_own, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        const ShardPlan& sp = e.c.shard_plan(level, world, rank);
        const size_t N = e.c.N;
        const size_t w = (size_t)2 * sp.nq_own * N;
        op_ks_shard_end(e, sp, dense(d_accP_all, (size_t)world * 2 * sp.cp * N), dense(d_acc_own, (size_t)2 * sp.n_own * N),
                        dense(d_addend_own ? d_addend_own : d_out_own, w), d_addend_own != nullptr, dense(d_out_own, w), batch);
    });
}


int lsb_rotate_hoisted(lsb_context_t h, int level, const uint64_t* d_ct, int n_rot, const uint64_t* const* d_glk, int key_level,
                       const uint64_t* gals, uint64_t* const* d_out, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        LSB_CHECK(n_rot >= 1 && d_glk && gals && d_out, "lsb: rotate_hoisted: bad arguments");
        const size_t w = e.c.ct_words(level, 2);
        std::vector<BatchPtr> outs(n_rot);
        for (int r = 0; r < n_rot; r++) outs[r] = dense(d_out[r], w);
        op_rotate_hoisted(e, level, dense(d_ct, w), n_rot, d_glk, key_level, gals, outs.data(), batch);
    });
}
int lsb_mac(lsb_context_t h, int level, int nterms, const uint64_t* const* d_cts, const uint64_t* const* d_pts, int ringt,
            const uint64_t* d_acc, uint64_t* d_out, int batch, void* stream) {
    return run_op(h, stream, [&](OpEnv& e) {
        LSB_CHECK(nterms >= 1 && d_cts && d_pts, "lsb: mac needs at least one term");
        const size_t cw = e.c.ct_words(level, 2), pw = ringt ? (size_t)e.c.N : e.c.ct_words(level, 1);
        std::vector<BatchPtr> cts(nterms), pts(nterms);
        for (int i = 0; i < nterms; i++) {
            cts[i] = dense(d_cts[i], cw);
            pts[i] = dense(d_pts[i], pw);
        }
        op_mac(e, level, nterms, cts.data(), pts.data(), ringt != 0, d_acc != nullptr, dense(d_acc ? d_acc : d_out, cw), dense(d_out, cw), batch);
    });
}

// ---- bootstrapping ----
struct lsb_bootstrap_st {
    lsb_context_t ctx;
    std::unique_ptr<DeviceBootstrap> b;
};
lsb_bootstrap_t lsb_bootstrap_create(lsb_context_t h, const int* ip, const double* dp) {
    LSB_API_BEGIN
    LSB_CHECK(h && h->c && ip && dp, "lsb: null argument");
    Context& c = *h->c;
    btp::Params P;
    P.logN = ip[0];
    P.log_slots = ip[1];
    P.q.assign(c.moduli.begin(), c.moduli.begin() + c.nq);
    P.p.assign(c.moduli.begin() + c.nq, c.moduli.begin() + c.nq + c.np);
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
    auto* o = new lsb_bootstrap_st();
    o->ctx = h;
    try {
        o->b.reset(new DeviceBootstrap(c, P));
    } catch (...) {
        delete o;
        throw;
    }
    return o;
    LSB_API_END(nullptr)
}
void lsb_bootstrap_destroy(lsb_bootstrap_t h) { delete h; }
int lsb_bootstrap_rotations(lsb_bootstrap_t h, int* steps, int capacity) {
    LSB_API_BEGIN
    LSB_CHECK(h && h->b, "lsb: null bootstrap handle");
    const std::vector<int> r = h->b->rotations();
    for (int i = 0; i < (int)r.size() && i < capacity; i++) steps[i] = r[i];
    return (int)r.size();
    LSB_API_END(-1)
}
int lsb_bootstrap_run(lsb_bootstrap_t h, const uint64_t* d_ct0, const uint64_t* d_rlk, int key_level, const uint64_t* gals,
                      const uint64_t* const* d_glk, int n_glk, const uint64_t* d_swk_dts, const uint64_t* d_swk_std, uint64_t* d_out,
                      int batch, void* stream, long long* stats5) {
    LSB_API_BEGIN
    LSB_CHECK(h && h->b && h->ctx && h->ctx->c, "lsb: null bootstrap handle");
    Context& c = *h->ctx->c;
    std::lock_guard<std::mutex> lk(h->ctx->mu);
    order_after_previous_stream(h->ctx, (cudaStream_t)stream);
    BootstrapKeys k;
    k.rlk = d_rlk;
    k.key_level = key_level;
    k.swk_dts = d_swk_dts;
    k.swk_std = d_swk_std;
    k.swk_dts_level = 0;
    for (int i = 0; i < n_glk; i++) {
        k.gal.push_back(gals[i]);
        k.glk.push_back(d_glk[i]);
    }
    BootstrapStats st;
    h->b->run(c.default_workspace(), (cudaStream_t)stream, k, bp_strided(d_ct0, (size_t)2 * c.N),
              bp_strided(d_out, (size_t)2 * (h->b->out_level() + 1) * c.N), batch, &st);
    if (stats5) {
        stats5[0] = st.rotations;
        stats5[1] = st.key_switches;
        stats5[2] = st.mults;
        stats5[3] = st.rescales;
        stats5[4] = st.plain_mults;
    }
    return 0;
    LSB_API_END(1)
}

}  // extern "C"
