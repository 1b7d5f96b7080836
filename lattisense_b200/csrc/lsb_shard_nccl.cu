// Host driver of the limb-sharded hybrid key-switch (the reference has no multi-GPU code at all,
// mega_ag_runners/gpu/gpu_wrapper.cu:141-391 is single-device): three local phases (lsb_shard.cu) around the path's only
// two exchange steps, issued as NCCL all-gathers over NVLink on the caller's stream -- all in C++ behind lsb200.h, so that
// a C/C++ host (one process per GPU) drives it without Python.  Batched over B ciphertexts; after a warm-up call the
// whole sequence (kernels + collectives) is captured once per argument set into a CUDA graph and replayed, because one
// un-batched key-switch is launch-latency bound (~25 launches).
//
// NCCL is resolved at run time (dlopen of libnccl.so.2: the copy PyTorch already loaded in a torchrun process, the system
// one otherwise), so the library has no link-time dependency on it; the communicator is created from a unique id that the
// caller distributes (lsb_nccl_unique_id on one rank, any broadcast -- torch.distributed, MPI, a file -- to the others).
#include <dlfcn.h>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>

#include "../../include/lsb200.h"
#include "lsb_ops.h"

using namespace lsb;

namespace {

// the slice of nccl.h this file needs (stable C ABI since NCCL 2.x)
typedef struct ncclComm* ncclComm_t;
typedef struct {
    char internal[128];
} ncclUniqueId;
typedef int ncclResult_t;
enum { ncclSuccess = 0 };
enum { ncclUint64 = 5 };  // ncclDataType_t: int8 0, uint8 1, int32 2, uint32 3, int64 4, uint64 5

struct Nccl {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

Nccl& nccl() {
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) {
            n.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (n.lib) break;
        }
        if (!n.lib) return;
        auto sym = [&](const char* s) { return dlsym(n.lib, s); };
        n.GetUniqueId = (decltype(n.GetUniqueId))sym("ncclGetUniqueId");
        n.CommInitRank = (decltype(n.CommInitRank))sym("ncclCommInitRank");
        n.CommDestroy = (decltype(n.CommDestroy))sym("ncclCommDestroy");
        n.AllGather = (decltype(n.AllGather))sym("ncclAllGather");
        n.GroupStart = (decltype(n.GroupStart))sym("ncclGroupStart");
        n.GroupEnd = (decltype(n.GroupEnd))sym("ncclGroupEnd");
        n.GetErrorString = (decltype(n.GetErrorString))sym("ncclGetErrorString");
    });
    if (!n.lib || !n.GetUniqueId || !n.CommInitRank || !n.AllGather || !n.GroupStart || !n.GroupEnd)
        throw std::runtime_error("lsb: NCCL (libnccl.so.2) could not be loaded: the limb-sharded key-switch driver needs it");
    return n;
}
#define LSB_NCCL(call)                                                                                                   \
    do {                                                                                                                 \
        ncclResult_t _r = (call);                                                                                        \
        if (_r != ncclSuccess)                                                                                           \
            throw std::runtime_error(std::string(#call) + ": " + (nccl().GetErrorString ? nccl().GetErrorString(_r) : "NCCL error")); \
    } while (0)

thread_local std::string g_err;

}  // namespace

struct lsb_context_st;  // lsb_capi.cu
Context& lsb_context_ref(lsb_context_t h);  // lsb_capi.cu

struct lsb_ks_shard_st {
    Context* c = nullptr;
    int level = 0, world = 1, rank = 0, max_batch = 1;
    ncclComm_t comm = nullptr;
    const ShardPlan* sp = nullptr;
    u64 *coeff_own = nullptr, *coeff_all = nullptr, *acc_own = nullptr, *accP_own = nullptr, *accP_all = nullptr;
    Workspace* ws = nullptr;
    std::unique_ptr<Workspace> ws_own;
    bool use_graph = true;
    struct Key {
        const void *c, *evk, *add, *out;
        int batch;
        cudaStream_t s;
        bool operator<(const Key& o) const { return std::tie(c, evk, add, out, batch, s) < std::tie(o.c, o.evk, o.add, o.out, o.batch, o.s); }
    };
    struct Cached {
        int seen = 0;
        cudaGraphExec_t exec = nullptr;
    };
    std::map<Key, Cached> graphs;
    long long all_gather_bytes = 0;  // per run, sent by this rank

    ~lsb_ks_shard_st() {
        if (c) cudaSetDevice(c->device);
        for (auto& kv : graphs)
            if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
        for (u64* p : {coeff_own, coeff_all, acc_own, accP_own, accP_all})
            if (p) cudaFree(p);
        if (comm && nccl().CommDestroy) nccl().CommDestroy(comm);
    }

    void enqueue(const u64* c_own, const u64* evk_own, const u64* addend_own, u64* out_own, int B, cudaStream_t s) {
        const size_t N = c->N;
        OpEnv e{*c, *ws, s, 0};
        const ShardPlan& p = *sp;
        const size_t w_c = (size_t)p.cq * N, w_ca = (size_t)world * p.cq * N, w_acc = (size_t)2 * std::max(p.n_own, 1) * N;
        const size_t w_p = (size_t)2 * p.cp * N, w_pa = (size_t)world * 2 * p.cp * N, w_in = (size_t)p.nq_own * N, w_out = (size_t)2 * p.nq_own * N;
        Nccl& n = nccl();
        op_ks_shard_begin(e, p, bp_strided(c_own, w_in), bp_strided(coeff_own, w_c), B);
        LSB_NCCL(n.GroupStart());
        for (int b = 0; b < B; b++) LSB_NCCL(n.AllGather(coeff_own + b * w_c, coeff_all + b * w_ca, w_c, ncclUint64, comm, s));
        LSB_NCCL(n.GroupEnd());
        op_ks_shard_inner(e, p, bp_strided(coeff_all, w_ca), bp_strided(c_own, w_in), evk_own, bp_strided(acc_own, w_acc), bp_strided(accP_own, w_p), B);
        LSB_NCCL(n.GroupStart());
        for (int b = 0; b < B; b++) LSB_NCCL(n.AllGather(accP_own + b * w_p, accP_all + b * w_pa, w_p, ncclUint64, comm, s));
        LSB_NCCL(n.GroupEnd());
        op_ks_shard_end(e, p, bp_strided(accP_all, w_pa), bp_strided(acc_own, w_acc), bp_strided(addend_own ? addend_own : out_own, w_out),
                        addend_own != nullptr, bp_strided(out_own, w_out), B);
        all_gather_bytes = (long long)B * (long long)(w_c + w_p) * 8;
    }
};

extern "C" {

int lsb_nccl_unique_id(void* id128) {
    g_err.clear();
    try {
        LSB_CHECK(id128, "lsb: null id buffer");
        ncclUniqueId id;
        LSB_NCCL(nccl().GetUniqueId(&id));
        std::memcpy(id128, &id, sizeof id);
        return 0;
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return 1;
    }
}
const char* lsb_ks_shard_last_error(void) { return g_err.c_str(); }

lsb_ks_shard_t lsb_ks_shard_create(lsb_context_t ctx, int level, int world, int rank, const void* nccl_unique_id128, int max_batch) {
    g_err.clear();
    try {
        LSB_CHECK(ctx && nccl_unique_id128 && world >= 1 && rank >= 0 && rank < world && max_batch >= 1, "lsb: bad arguments to lsb_ks_shard_create");
        Context& c = lsb_context_ref(ctx);
        LSB_CUDA(cudaSetDevice(c.device));
        std::unique_ptr<lsb_ks_shard_st> h(new lsb_ks_shard_st());
        h->c = &c;
        h->level = level;
        h->world = world;
        h->rank = rank;
        h->max_batch = max_batch;
        h->sp = &c.shard_plan(level, world, rank);
        h->ws_own.reset(new Workspace(c.device));
        h->ws = h->ws_own.get();
        if (const char* g = getenv("LSB_KS_SHARD_GRAPH")) h->use_graph = atoi(g) != 0;
        ncclUniqueId id;
        std::memcpy(&id, nccl_unique_id128, sizeof id);
        LSB_NCCL(nccl().CommInitRank(&h->comm, world, id, rank));
        const ShardPlan& p = *h->sp;
        const size_t N = c.N, B = (size_t)max_batch;
        auto alloc = [&](size_t words) {
            u64* d = nullptr;
            LSB_CUDA(cudaMalloc((void**)&d, std::max<size_t>(words, 1) * sizeof(u64)));
            LSB_CUDA(cudaMemset(d, 0, std::max<size_t>(words, 1) * sizeof(u64)));
            return d;
        };
        h->coeff_own = alloc(B * p.cq * N);
        h->coeff_all = alloc(B * world * p.cq * N);
        h->acc_own = alloc(B * 2 * std::max(p.n_own, 1) * N);
        h->accP_own = alloc(B * 2 * p.cp * N);
        h->accP_all = alloc(B * world * 2 * p.cp * N);
        return h.release();
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return nullptr;
    }
}
void lsb_ks_shard_destroy(lsb_ks_shard_t h) { delete h; }

int lsb_ks_shard_run(lsb_ks_shard_t h, const uint64_t* d_c_own_ntt, const uint64_t* d_evk_own, const uint64_t* d_addend_own,
                     uint64_t* d_out_own, int batch, void* stream) {
    g_err.clear();
    try {
        LSB_CHECK(h && h->c && d_c_own_ntt && d_evk_own && d_out_own, "lsb: null argument to lsb_ks_shard_run");
        LSB_CHECK(batch >= 1 && batch <= h->max_batch, "lsb: batch exceeds the driver's max_batch");
        LSB_CUDA(cudaSetDevice(h->c->device));
        cudaStream_t s = (cudaStream_t)stream;
        if (!h->use_graph || s == nullptr) {  // the legacy default stream cannot be captured
            h->enqueue(d_c_own_ntt, d_evk_own, d_addend_own, d_out_own, batch, s);
            return 0;
        }
        lsb_ks_shard_st::Key key{d_c_own_ntt, d_evk_own, d_addend_own, d_out_own, batch, s};
        lsb_ks_shard_st::Cached& cg = h->graphs[key];
        if (cg.exec) {
            LSB_CUDA(cudaGraphLaunch(cg.exec, s));
            return 0;
        }
        if (cg.seen++ == 0) {  // first sight: run eagerly (grows the workspace, builds the cached tables, warms NCCL up)
            h->enqueue(d_c_own_ntt, d_evk_own, d_addend_own, d_out_own, batch, s);
            return 0;
        }
        cudaGraph_t graph = nullptr;
        LSB_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        try {
            h->enqueue(d_c_own_ntt, d_evk_own, d_addend_own, d_out_own, batch, s);
        } catch (...) {
            cudaStreamEndCapture(s, &graph);
            if (graph) cudaGraphDestroy(graph);
            throw;
        }
        LSB_CUDA(cudaStreamEndCapture(s, &graph));
        LSB_CUDA(cudaGraphInstantiate(&cg.exec, graph, 0));
        cudaGraphDestroy(graph);
        LSB_CUDA(cudaGraphLaunch(cg.exec, s));
        return 0;
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return 1;
    }
}

long long lsb_ks_shard_all_gather_bytes(lsb_ks_shard_t h) { return h ? h->all_gather_bytes : 0; }

}  // extern "C"
