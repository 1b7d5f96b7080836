// libgpu_mega_ag_runner.so — the mega_ag GPU executor behind create_/bind_/run_/release_fhe_gpu_task.
//
// Replaces mega_ag_runners/gpu/gpu_wrapper.cu (FheGpuTask, _run_mega_ag_impl :141-391), the per-node HEonGPU
// binder mega_ag_executors_gpu.cu:463 and the load/store executors of gpu_abi_bridge_executors.h:213-363.
// Differences by design (B200-first, same observable behaviour):
//   * the RNS context / NTT tables are built once per (task, device) and reused (the reference rebuilds them
//     on every run, gpu_wrapper.cu:152-155);
//   * ready GPU nodes with the same signature (operation, level, key) are executed as ONE batched launch train
//     over a device-pointer table instead of one launch train per node on 2 streams;
//   * mult -> relin (-> rescale) chains whose intermediates have no other consumer run as one fused op;
//   * LOAD copies an object's limbs into a pinned staging ring and issues one H2D copy per object (the
//     reference issues one pageable cudaMemcpyAsync per limb); STORE is one D2H per object;
//   * independent components of the DAG can be sharded over several GPUs (lsb_gpu_task_set_devices);
//   * the first failing node cancels the run and run_fhe_gpu_task reports it (the reference swallows
//     exceptions of CPU-pool nodes, cpu_task_utils.h:450-459).
#include <any>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <queue>
#include <thread>
#include <unordered_map>
#include <unordered_set>

#include "lattisense_gpu_runner.h"
#include "lattisense_mega_ag.hpp"
#include "lsb_ops.h"
#include "lsb_bootstrap_dev.h"
#include "lsr_graph.h"
#include "lsr_abi_arena.h"

namespace lsr {

using lsb::BatchPtr;
using lsb::u64;

void lsr_stream_copy(void* dst, const void* src, size_t bytes);  // lsr_hostcopy.cpp

// ---------------------------------------------------------------------------------------------
// host worker pool (EXPORT / IMPORT / custom nodes, staging copies)
// ---------------------------------------------------------------------------------------------
class WorkerPool {
public:
    explicit WorkerPool(int n) {
        for (int i = 0; i < n; i++) threads_.emplace_back([this] { loop(); });
    }
    ~WorkerPool() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : threads_) t.join();
    }
    void submit(int priority, std::function<void()> fn) {
        {
            std::lock_guard<std::mutex> lk(mu_);
            q_.push(Item{priority, seq_++, std::move(fn)});
        }
        cv_.notify_one();
    }
    int size() const { return (int)threads_.size(); }

private:
    struct Item {
        int pri;
        uint64_t seq;
        std::function<void()> fn;
        bool operator<(const Item& o) const { return pri != o.pri ? pri < o.pri : seq > o.seq; }
    };
    void loop() {
        for (;;) {
            std::function<void()> fn;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [this] { return stop_ || !q_.empty(); });
                if (q_.empty()) return;
                fn = std::move(const_cast<Item&>(q_.top()).fn);
                q_.pop();
            }
            fn();
        }
    }
    std::mutex mu_;
    std::condition_variable cv_;
    std::priority_queue<Item> q_;
    std::vector<std::thread> threads_;
    uint64_t seq_ = 0;
    bool stop_ = false;
};

// ---------------------------------------------------------------------------------------------
// per-device runtime: RNS context, streams, pinned staging rings
// ---------------------------------------------------------------------------------------------
struct PinnedRing {
    // Ring allocator over one pinned buffer, shared by the worker threads.  acquire() hands out a region
    // (blocking while older regions that overlap it are still in use); the owner fills it / has it
    // filled, then commit()s it with the CUDA event after which it may be reused.
    uint8_t* base = nullptr;
    size_t cap = 0, head = 0;
    struct Busy {
        uint64_t ticket;
        size_t begin, end;
        cudaEvent_t ev;  // nullptr until committed
        bool done;       // released by ticket (regions lent out as zero-copy C structs)
    };
    std::deque<Busy> busy;
    std::vector<cudaEvent_t> free_events;
    std::mutex mu;
    std::condition_variable cv;
    uint64_t next_ticket = 1;

    void init(size_t bytes) {
        LSB_CUDA(cudaHostAlloc((void**)&base, bytes, cudaHostAllocDefault));
        cap = bytes;
        lsr::register_pinned_range(base, bytes);  // limbs lent to ABI structs must never reach free() (lsr_abi.cpp)
    }
    void destroy() {
        for (auto& b : busy)
            if (b.ev) cudaEventDestroy(b.ev);
        for (auto e : free_events) cudaEventDestroy(e);
        busy.clear();
        free_events.clear();
        if (base) {
            lsr::unregister_pinned_range(base);
            cudaFreeHost(base);
        }
        base = nullptr;
    }
    cudaEvent_t new_event() {
        std::lock_guard<std::mutex> lk(mu);
        if (!free_events.empty()) {
            cudaEvent_t e = free_events.back();
            free_events.pop_back();
            return e;
        }
        cudaEvent_t e;
        LSB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        return e;
    }
    void drop_ticket(uint64_t ticket) {  // mu held
        for (auto it = busy.begin(); it != busy.end(); ++it)
            if (it->ticket == ticket) {
                if (it->ev) free_events.push_back(it->ev);
                busy.erase(it);
                return;
            }
    }
    uint8_t* acquire(size_t bytes, uint64_t* ticket) {
        bytes = (bytes + 255) & ~(size_t)255;
        LSB_CHECK(bytes <= cap, "lsb runner: object larger than the pinned staging ring (raise LSB_STAGING_MB)");
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            // drop finished regions from the front
            while (!busy.empty() && (busy.front().done || (busy.front().ev && cudaEventQuery(busy.front().ev) == cudaSuccess))) {
                if (busy.front().ev) free_events.push_back(busy.front().ev);
                busy.pop_front();
            }
            size_t at = head + bytes > cap ? 0 : head;
            const Busy* clash = nullptr;
            for (auto& b : busy)
                if (at < b.end && b.begin < at + bytes) {
                    clash = &b;
                    break;
                }
            if (!clash) {
                head = at + bytes;
                *ticket = next_ticket++;
                busy.push_back(Busy{*ticket, at, at + bytes, nullptr, false});
                return base + at;
            }
            const uint64_t ct = clash->ticket;
            if (clash->done) {
                drop_ticket(ct);
            } else if (clash->ev) {
                cudaEvent_t ev = clash->ev;
                lk.unlock();
                LSB_CUDA(cudaEventSynchronize(ev));
                lk.lock();
                // the entry may not be at the front (regions finish out of order): drop it by ticket unless another
                // thread already did and the event was recycled for a newer region
                for (auto& b2 : busy)
                    if (b2.ticket == ct && b2.ev == ev) {
                        drop_ticket(ct);
                        break;
                    }
            } else {
                cv.wait_for(lk, std::chrono::milliseconds(1));
            }
        }
    }
    void commit(uint64_t ticket, cudaEvent_t ev) {
        {
            std::lock_guard<std::mutex> lk(mu);
            for (auto& b : busy)
                if (b.ticket == ticket) b.ev = ev;
        }
        cv.notify_all();
    }
    // the region is free again regardless of any event (its owner is done with the memory)
    void release(uint64_t ticket) {
        {
            std::lock_guard<std::mutex> lk(mu);
            for (auto& b : busy)
                if (b.ticket == ticket) b.done = true;
        }
        cv.notify_all();
    }
};

struct DeviceRuntime {
    int device = 0;
    std::unique_ptr<lsb::Context> ctx;
    std::unique_ptr<lsb::Workspace> ws;
    cudaStream_t compute = nullptr, h2d = nullptr, d2h = nullptr;
    // extra upload streams: object copies alternate over them so that two copy engines keep the link busy across the
    // per-transfer gaps of 1 MiB objects (LSB_H2D_STREAMS; default 1: measured no gain on cfg2, the path is host-DRAM bound)
    std::vector<cudaStream_t> h2d_extra;
    std::atomic<unsigned> h2d_rr{0};
    cudaStream_t next_h2d() {
        if (h2d_extra.empty()) return h2d;
        const unsigned i = h2d_rr.fetch_add(1) % (unsigned)(h2d_extra.size() + 1);
        return i == 0 ? h2d : h2d_extra[i - 1];
    }
    std::mutex mu;  // serialises enqueues on the copy streams (worker threads issue copies)
    PinnedRing up, down;
    // device pointer tables of batched launches: pinned host mirror + device arena, both rings
    u64** tab_host = nullptr;
    u64** tab_dev = nullptr;
    size_t tab_cap = 0, tab_head = 0;
    cudaEvent_t tab_wrap_ev = nullptr;
    // bootstrapping precompute (encoded CoeffsToSlots / SlotsToCoeffs matrices), built at the first bootstrap node and
    // kept with the task: the analogue of generate_bootstrapping_params_v2 in init_gpu_context (gpu_wrapper.cu:106-116)
    std::unique_ptr<lsb::DeviceBootstrap> btp;

    ~DeviceRuntime() {
        cudaSetDevice(device);
        if (compute) cudaStreamSynchronize(compute);
        if (h2d) cudaStreamSynchronize(h2d);
        for (auto st : h2d_extra) cudaStreamSynchronize(st);
        if (d2h) cudaStreamSynchronize(d2h);
        up.destroy();
        down.destroy();
        if (tab_host) cudaFreeHost(tab_host);
        if (tab_dev) cudaFree(tab_dev);
        if (tab_wrap_ev) cudaEventDestroy(tab_wrap_ev);
        btp.reset();
        ws.reset();
        ctx.reset();
        if (compute) cudaStreamDestroy(compute);
        if (h2d) cudaStreamDestroy(h2d);
        for (auto st : h2d_extra) cudaStreamDestroy(st);
        if (d2h) cudaStreamDestroy(d2h);
        // the release threshold of the stream-ordered pool is "never" while tasks run (blocks are recycled across runs);
        // a released task hands its cached blocks back, so that a process that goes on to other work on this device does
        // not allocate out of a pool fragmented into this task's object sizes
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
    }
};

static size_t env_mb(const char* name, size_t dflt) {
    const char* e = getenv(name);
    return e ? (size_t)atol(e) : dflt;
}

static std::unique_ptr<DeviceRuntime> make_device_runtime(int device, const TaskGraph& g) {
    int ndev = 0;
    cudaError_t err = cudaGetDeviceCount(&ndev);
    if (err != cudaSuccess || ndev == 0)
        throw std::runtime_error(std::string("lsb runner: no CUDA device available (") + cudaGetErrorString(err) +
                                 "); this executor has no CPU fallback");
    LSB_CHECK(device >= 0 && device < ndev, "lsb runner: gpu_device out of range");
    std::unique_ptr<DeviceRuntime> rt(new DeviceRuntime());
    rt->device = device;
    LSB_CUDA(cudaSetDevice(device));
    int logn = 0;
    while ((1 << logn) < g.param.n) logn++;
    LSB_CHECK((1 << logn) == g.param.n, "lsb runner: ring degree is not a power of two");
    const int nq = std::min((int)g.param.q.size(), g.param.max_level + 1);
    rt->ctx.reset(new lsb::Context(device, g.algo == ALGO_BFV ? lsb::SCHEME_BFV : lsb::SCHEME_CKKS, logn, g.param.q.data(),
                                   nq, g.param.p.data(), (int)g.param.p.size(), g.algo == ALGO_BFV ? g.param.t : 0));
    rt->ws.reset(new lsb::Workspace(device));
    LSB_CUDA(cudaStreamCreateWithFlags(&rt->compute, cudaStreamNonBlocking));
    LSB_CUDA(cudaStreamCreateWithFlags(&rt->h2d, cudaStreamNonBlocking));
    for (size_t i = 1; i < std::max<size_t>(1, env_mb("LSB_H2D_STREAMS", 1)); i++) {
        cudaStream_t st;
        LSB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        rt->h2d_extra.push_back(st);
    }
    LSB_CUDA(cudaStreamCreateWithFlags(&rt->d2h, cudaStreamNonBlocking));
    rt->up.init(env_mb("LSB_STAGING_MB", 512) << 20);
    rt->down.init(env_mb("LSB_STAGING_MB", 512) << 20);
    rt->tab_cap = 1 << 20;
    LSB_CUDA(cudaHostAlloc((void**)&rt->tab_host, rt->tab_cap * sizeof(u64*), cudaHostAllocDefault));
    LSB_CUDA(cudaMalloc((void**)&rt->tab_dev, rt->tab_cap * sizeof(u64*)));
    LSB_CUDA(cudaEventCreateWithFlags(&rt->tab_wrap_ev, cudaEventDisableTiming));
    // keep freed blocks of the stream-ordered allocator cached across runs
    cudaMemPool_t pool;
    LSB_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t thresh = UINT64_MAX;
    LSB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh));
    return rt;
}

// ---------------------------------------------------------------------------------------------
// device-resident objects (what LOAD produces, GPU nodes consume/produce and STORE reads)
// ---------------------------------------------------------------------------------------------
struct DeviceObject {
    DeviceRuntime* rt = nullptr;
    u64* ptr = nullptr;
    size_t words = 0;
    int level = 0;
    int npoly = 0;      // ciphertext: degree+1; plaintext: 1; key: 0
    int limbs = 0;      // limbs per polynomial
    cudaEvent_t ready = nullptr;  // set when produced on a copy stream (compute stream must wait on it)

    ~DeviceObject() {
        if (!rt) return;
        cudaSetDevice(rt->device);
        // all consumers were enqueued on the compute stream (or finished, for stores) before the last
        // reference is dropped, so a free ordered on that stream is safe -- once that stream also waits for the copy
        // that PRODUCED the object (a LOADed object of a failed / cancelled run may never have had a consumer)
        if (ptr && ready) cudaStreamWaitEvent(rt->compute, ready, 0);
        if (ptr) cudaFreeAsync(ptr, rt->compute);
        if (ready) cudaEventDestroy(ready);
    }
};
using DevObjPtr = std::shared_ptr<DeviceObject>;

static DevObjPtr alloc_device_object(DeviceRuntime* rt, size_t words, cudaStream_t s) {
    DevObjPtr o = std::make_shared<DeviceObject>();
    o->rt = rt;
    o->words = words;
    LSB_CUDA(cudaSetDevice(rt->device));
    LSB_CUDA(cudaMallocAsync((void**)&o->ptr, words * sizeof(u64), s));
    return o;
}


// A datum living on the device(s): keys are replicated on every device that consumes them, anything
// else lives on the one device its DAG component was assigned to.
struct DeviceDatum {
    std::vector<DevObjPtr> on;  // indexed by device slot
};
using DatumPtr = std::shared_ptr<DeviceDatum>;

// ---------------------------------------------------------------------------------------------
// execution plan: units (one node, or a fused mult -> relin [-> rescale] chain), device assignment
// ---------------------------------------------------------------------------------------------
enum class UnitKind { HOST, LOAD, STORE, GPU };

struct Unit {
    int id = 0;
    UnitKind kind = UnitKind::HOST;
    std::vector<ComputeNode*> nodes;       // 1 node, or the chain in execution order
    std::vector<DatumNode*> ext_inputs;    // distinct data read from outside the unit
    DatumNode* output = nullptr;
    int priority = 0;
    int slot = 0;                          // device slot; -1: every device (key load)
    int component = -1;
    int order_key = 0;                     // LOAD: id of the first GPU unit that consumes the object (-1: key)
    // re-associated sum (nodes = absorbed products, inner additions, root addition last): out = sum ct [* pt]
    std::vector<std::pair<DatumNode*, DatumNode*>> sum_terms;  // (ciphertext, plaintext or nullptr)
};

struct Plan {
    std::vector<Unit> units;
    std::unordered_map<NodeIndex, std::vector<int>> consumers;  // datum -> units reading it
    std::unordered_map<NodeIndex, int> producer;                // datum -> unit writing it
    int n_components = 0;
    std::vector<double> component_weight;
    long long fused_nodes = 0;
    size_t total_nodes = 0;
};

static OperationType op_of(const ComputeNode* c) {
    return c->fhe_prop.has_value() ? c->fhe_prop->op_type : OperationType::UNKNOWN;
}
static bool is_device_op(const ComputeNode* c) {
    if (c->custom_prop.has_value() || !c->fhe_prop.has_value()) return false;
    switch (c->fhe_prop->op_type) {
        case OperationType::EXPORT_TO_ABI:
        case OperationType::IMPORT_FROM_ABI:
        case OperationType::LOAD_TO_BACKEND:
        case OperationType::STORE_FROM_BACKEND: return false;
        default: return true;
    }
}
// the single consumer of `d` if it is a device op of type `want` and `d` is a pure intermediate
static ComputeNode* sole_consumer(const DatumNode* d, OperationType want) {
    if (d->is_output || d->is_input || d->successors.size() != 1) return nullptr;
    ComputeNode* c = d->successors[0];
    if (!is_device_op(c) || op_of(c) != want) return nullptr;
    return c->input_nodes[0] == d ? c : nullptr;
}

// Keys and plaintexts are read-only inputs that several otherwise independent sub-graphs may share (cfg3 x B: one set of
// polynomial coefficients for every evaluation): they do not join DAG components, each is loaded onto every device
// that runs one of its consumers.
static bool replicated_datum(const DatumNode* d) {
    return d->fhe_prop.has_value() && (is_key_type(d->datum_type) || d->datum_type == TYPE_PLAINTEXT);
}

static Plan build_plan(TaskGraph& g, bool fuse, bool fuse_host) {
    Plan p;
    p.total_nodes = g.computes.size();
    std::vector<NodeIndex> order;
    for (auto& kv : g.computes) order.push_back(kv.first);
    std::sort(order.begin(), order.end());
    std::unordered_set<ComputeNode*> absorbed;
    // pass 1: fused chains start at a ct x ct multiply
    std::unordered_map<ComputeNode*, std::vector<ComputeNode*>> chains;
    if (fuse) {
        for (NodeIndex ci : order) {
            ComputeNode* m = &g.computes.at(ci);
            if (!is_device_op(m) || op_of(m) != OperationType::MULTIPLY || m->input_nodes.empty() || m->input_nodes.size() > 2) continue;
            bool ct_ct = true;  // one input = square (mega_ag_executors_gpu.cu:178-187)
            for (DatumNode* in : m->input_nodes)
                ct_ct = ct_ct && in->fhe_prop.has_value() && in->datum_type == TYPE_CIPHERTEXT && in->fhe_prop->degree == 1;
            if (!ct_ct) continue;
            ComputeNode* r = sole_consumer(m->output_nodes[0], OperationType::RELINEARIZE);
            if (!r) continue;
            std::vector<ComputeNode*> chain{m, r};
            if (ComputeNode* s = sole_consumer(r->output_nodes[0], OperationType::RESCALE)) chain.push_back(s);
            for (size_t i = 1; i < chain.size(); i++) absorbed.insert(chain[i]);
            chains[m] = chain;
        }
    }
    // pass 1a': sums.  A tree of ct + ct additions whose inner results have no other reader becomes ONE n-ary sum, and a
    // CKKS ct x pt product read only by that sum becomes a multiply-accumulate term of it.  The packed convolution of
    // examples/benchmark_convolution accumulates 12 800 products through a chain of 12 800 dependent additions
    // (benchmark_convolution.py:143-156): node by node that is 12 800 scheduling waves, re-associated it is one pass over
    // the terms.  Modular addition is associative and commutative, so every limb is unchanged.
    std::unordered_map<ComputeNode*, std::vector<std::pair<DatumNode*, DatumNode*>>> sums;
    if (fuse) {
        auto is_ct = [](const DatumNode* d) { return d->fhe_prop.has_value() && d->datum_type == TYPE_CIPHERTEXT; };
        auto is_sum_add = [&](const ComputeNode* c) {
            if (!is_device_op(c) || op_of(c) != OperationType::ADD || c->input_nodes.size() != 2 || c->output_nodes.size() != 1) return false;
            const DatumNode *a = c->input_nodes[0], *b = c->input_nodes[1], *o = c->output_nodes[0];
            return is_ct(a) && is_ct(b) && is_ct(o) && a->fhe_prop->level == b->fhe_prop->level && a->fhe_prop->degree == b->fhe_prop->degree &&
                   o->fhe_prop->level == a->fhe_prop->level && o->fhe_prop->degree == a->fhe_prop->degree;
        };
        auto pure_single = [](const DatumNode* d) { return !d->is_output && !d->is_input && d->successors.size() == 1 && d->predecessors.size() == 1; };
        for (NodeIndex ci : order) {
            ComputeNode* root = &g.computes.at(ci);
            if (!is_sum_add(root) || absorbed.count(root)) continue;
            const DatumNode* ro = root->output_nodes[0];
            if (pure_single(ro) && is_sum_add(ro->successors[0])) continue;  // an inner node of somebody else's sum
            std::vector<ComputeNode*> inner, prods;
            std::vector<std::pair<DatumNode*, DatumNode*>> mac, plain;
            std::vector<DatumNode*> stack{root->input_nodes[1], root->input_nodes[0]};
            while (!stack.empty()) {
                DatumNode* d = stack.back();
                stack.pop_back();
                ComputeNode* pn = pure_single(d) ? d->predecessors[0] : nullptr;
                if (pn && !absorbed.count(pn) && is_sum_add(pn)) {
                    inner.push_back(pn);
                    stack.push_back(pn->input_nodes[1]);
                    stack.push_back(pn->input_nodes[0]);
                    continue;
                }
                if (pn && !absorbed.count(pn) && !chains.count(pn) && g.algo == ALGO_CKKS && is_device_op(pn) && op_of(pn) == OperationType::MULTIPLY &&
                    pn->input_nodes.size() == 2 && pn->output_nodes.size() == 1) {
                    DatumNode *x = pn->input_nodes[0], *y = pn->input_nodes[1];
                    if (!is_ct(x)) std::swap(x, y);
                    const bool plain_pt = y->fhe_prop.has_value() && y->datum_type == TYPE_PLAINTEXT &&
                                          !(y->fhe_prop->p.has_value() && y->fhe_prop->p->is_ringt);
                    if (is_ct(x) && plain_pt && x->fhe_prop->level == d->fhe_prop->level && x->fhe_prop->degree == d->fhe_prop->degree &&
                        y->fhe_prop->level == d->fhe_prop->level) {
                        prods.push_back(pn);
                        mac.emplace_back(x, y);
                        continue;
                    }
                }
                plain.emplace_back(d, nullptr);
            }
            if (inner.size() + prods.size() == 0) continue;  // a lone addition of two finished ciphertexts
            std::vector<ComputeNode*> nodes = prods;
            nodes.insert(nodes.end(), inner.rbegin(), inner.rend());
            nodes.push_back(root);
            for (size_t i = 0; i + 1 < nodes.size(); i++) absorbed.insert(nodes[i]);
            chains[root] = nodes;
            mac.insert(mac.end(), plain.begin(), plain.end());
            sums[root] = std::move(mac);
        }
    }
    // pass 1b: host chains.  EXPORT_TO_ABI -> LOAD_TO_BACKEND and STORE_FROM_BACKEND -> IMPORT_FROM_ABI run back to
    // back on one worker thread: the limbs EXPORT just wrote are still in that core's cache when LOAD stages them,
    // and IMPORT reads the stored object straight out of the pinned staging region (no intermediate copy).
    if (fuse_host) {
        for (NodeIndex ci : order) {
            ComputeNode* a = &g.computes.at(ci);
            if (a->custom_prop.has_value() || !a->fhe_prop.has_value() || a->output_nodes.size() != 1) continue;
            const OperationType op = op_of(a);
            if (op != OperationType::EXPORT_TO_ABI && op != OperationType::STORE_FROM_BACKEND) continue;
            const DatumNode* d = a->output_nodes[0];
            if (d->is_output || d->is_input || d->successors.size() != 1) continue;
            ComputeNode* b = d->successors[0];
            if (b->custom_prop.has_value() || !b->fhe_prop.has_value() || b->input_nodes.size() != 1 || absorbed.count(b)) continue;
            const OperationType want = op == OperationType::EXPORT_TO_ABI ? OperationType::LOAD_TO_BACKEND : OperationType::IMPORT_FROM_ABI;
            if (op_of(b) != want) continue;
            chains[a] = {a, b};
            absorbed.insert(b);
        }
    }
    for (NodeIndex ci : order) {
        ComputeNode* c = &g.computes.at(ci);
        if (absorbed.count(c)) continue;
        Unit u;
        u.id = (int)p.units.size();
        auto it = chains.find(c);
        if (it != chains.end()) {
            u.nodes = it->second;
            if (is_device_op(c)) p.fused_nodes += (long long)u.nodes.size() - 1;
            auto st = sums.find(c);
            if (st != sums.end()) u.sum_terms = st->second;
        } else {
            u.nodes = {c};
        }
        const OperationType op = op_of(u.nodes.back()) == OperationType::LOAD_TO_BACKEND ? OperationType::LOAD_TO_BACKEND : op_of(c);
        if (op == OperationType::LOAD_TO_BACKEND) u.kind = UnitKind::LOAD;
        else if (op == OperationType::STORE_FROM_BACKEND) u.kind = UnitKind::STORE;
        else if (is_device_op(c)) u.kind = UnitKind::GPU;
        else u.kind = UnitKind::HOST;
        std::unordered_set<const DatumNode*> internal;
        for (ComputeNode* n : u.nodes)
            for (DatumNode* o : n->output_nodes) internal.insert(o);
        for (ComputeNode* n : u.nodes) {
            u.priority = std::max(u.priority, n->priority);
            for (DatumNode* in : n->input_nodes) {
                if (internal.count(in)) continue;
                if (std::find(u.ext_inputs.begin(), u.ext_inputs.end(), in) == u.ext_inputs.end()) u.ext_inputs.push_back(in);
            }
        }
        ComputeNode* last = u.nodes.back();
        LSB_CHECK(last->output_nodes.size() == 1, "mega_ag: compute node '" + last->id + "' must have exactly one output");
        u.output = last->output_nodes[0];
        p.units.push_back(std::move(u));
    }
    for (Unit& u : p.units) {
        for (DatumNode* in : u.ext_inputs) p.consumers[in->index].push_back(u.id);
        p.producer[u.output->index] = u.id;
    }
    for (Unit& u : p.units) {
        if (u.kind != UnitKind::LOAD) continue;
        const bool key = u.output->fhe_prop.has_value() && is_key_type(u.output->datum_type);
        u.order_key = key ? -1 : (int)p.units.size();
        auto it = p.consumers.find(u.output->index);
        if (!key && it != p.consumers.end())
            for (int v : it->second) u.order_key = std::min(u.order_key, v);
    }
    // components: units joined through ciphertext data (keys and plaintexts are replicated instead)
    std::vector<int> parent(p.units.size());
    for (size_t i = 0; i < parent.size(); i++) parent[i] = (int)i;
    std::function<int(int)> find = [&](int x) {
        while (parent[x] != x) x = parent[x] = parent[parent[x]];
        return x;
    };
    for (Unit& u : p.units) {
        if (replicated_datum(u.output)) continue;
        auto it = p.consumers.find(u.output->index);
        if (it == p.consumers.end()) continue;
        for (int v : it->second) parent[find(v)] = find(u.id);
    }
    std::unordered_map<int, int> comp_id;
    for (Unit& u : p.units) {
        const bool key_chain = replicated_datum(u.output);
        if (key_chain) {
            u.component = -1;
            u.slot = -1;
            continue;
        }
        int r = find(u.id);
        auto it = comp_id.find(r);
        if (it == comp_id.end()) it = comp_id.emplace(r, (int)comp_id.size()).first;
        u.component = it->second;
    }
    p.n_components = (int)comp_id.size();
    p.component_weight.assign(p.n_components, 0.0);
    for (Unit& u : p.units) {
        if (u.component < 0) continue;
        double w = 0.05;
        if (u.kind == UnitKind::GPU) {
            w = 0.2;
            for (ComputeNode* n : u.nodes) {
                const OperationType op = op_of(n);
                if (op == OperationType::RELINEARIZE || op == OperationType::ROTATE_COL || op == OperationType::ROTATE_ROW) w += 4.0;
                else if (op == OperationType::MULTIPLY || op == OperationType::RESCALE) w += 1.0;
                else if (op == OperationType::BOOTSTRAP) w += 2000.0;
            }
        } else if (u.kind == UnitKind::LOAD || u.kind == UnitKind::STORE) {
            w = 0.5;
        }
        p.component_weight[u.component] += w;
    }
    return p;
}

// longest-processing-time-first assignment of components to device slots
static std::vector<int> assign_components(const Plan& p, int n_slots) {
    std::vector<int> slot_of(p.n_components, 0);
    if (n_slots <= 1) return slot_of;
    std::vector<int> idx(p.n_components);
    for (int i = 0; i < p.n_components; i++) idx[i] = i;
    std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return p.component_weight[a] > p.component_weight[b]; });
    std::vector<double> load(n_slots, 0.0);
    for (int c : idx) {
        int best = 0;
        for (int s = 1; s < n_slots; s++)
            if (load[s] < load[best]) best = s;
        slot_of[c] = best;
        load[best] += p.component_weight[c];
    }
    return slot_of;
}

// How many of the `n_slots` devices a run should open.  Every extra device costs a replica of the keys it needs (one H2D
// each), a dispatcher and shorter launch trains; measured on B200s with cfg3 x 64 (64 components, total weight ~3.8 k):
// 21.9 / 15.3 / 11.7 ms on 1 / 2 / 4 devices but 15.5 ms on 8.  A device is therefore opened only per
// LSB_MIN_WEIGHT_PER_DEVICE (default 800, about 200 key switches at n = 16384; scaled with the ring degree) of
// component weight, and never for more devices than there are components.  The first `result` devices of the list are used.
static int worthwhile_slots(const Plan& p, int n_slots, double ring_n) {
    if (n_slots <= 1) return std::max(1, n_slots);
    const char* env = getenv("LSB_MIN_WEIGHT_PER_DEVICE");  // read per run: tests turn the cap off with 0
    const double min_w = env ? std::max(0.0, atof(env)) : 800.0;
    double total = 0.0;
    for (double w : p.component_weight) total += w;
    total *= std::max(1.0, ring_n / 16384.0);
    int use = min_w > 0.0 ? (int)(total / min_w) : n_slots;
    use = std::min(use, p.n_components);
    return std::max(1, std::min(use, n_slots));
}

// ---------------------------------------------------------------------------------------------
// the task object behind fhe_task_handle
// ---------------------------------------------------------------------------------------------
struct GpuTask {
    std::unique_ptr<TaskGraph> graph;
    Plan plan;
    // by (CUDA device index, ordinal of the slot on that device), built on first use: lsb_gpu_task_set_devices may name a
    // device twice -- two independent slots (streams, dispatcher, staging rings, tables) sharing one GPU
    std::map<std::pair<int, int>, std::unique_ptr<DeviceRuntime>> runtimes;
    std::vector<int> devices;                                // lsb_gpu_task_set_devices; empty: run()'s gpu_device
    FrontendContextFactory ctx_factory;
    std::mutex run_mu;  // one run at a time per handle (like the reference: FheTask::input_args is reused)
    long long stat_launches = 0, stat_groups = 0, stat_fused = 0;
    // transfer accounting of the last run: bytes copied host->device / device->host, of which read straight out of the
    // pinned ring the EXPORT executor wrote into (no staging copy), and key bytes NOT uploaded thanks to the key cache
    std::atomic<long long> stat_h2d{0}, stat_d2h{0}, stat_zero_copy{0}, stat_key_cached{0};
    // host-side time of the last run, summed over worker threads (ns): EXPORT executors, staging copies, waiting for ring
    // space, waiting for a result's D2H, IMPORT executors; and the wall time of the run
    std::atomic<long long> ns_export{0}, ns_stage{0}, ns_ring_wait{0}, ns_d2h_wait{0}, ns_import{0}, ns_wall{0};
    // device-resident key cache across run() calls (the reference re-uploads every key on every run,
    // gpu_abi_bridge_executors.h:213-310): opt-in (lsb_gpu_task_set_key_cache / LSB_KEY_CACHE=1) because it assumes that
    // a key HANDLE that is passed again still holds the same key; lsb_gpu_task_invalidate_keys drops it
    bool key_cache_on = false;
    std::mutex key_mu;
    std::map<std::pair<const void*, NodeIndex>, std::any> key_cache;  // (caller's key handle, LOAD output datum) -> DatumPtr

    DeviceRuntime* runtime(int device, int ordinal = 0) {
        const std::pair<int, int> key(device, ordinal);
        auto it = runtimes.find(key);
        if (it == runtimes.end()) it = runtimes.emplace(key, make_device_runtime(device, *graph)).first;
        return it->second.get();
    }
};

static thread_local std::string g_runner_error;
static inline long long now_ns() {
    return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---------------------------------------------------------------------------------------------
// one run
// ---------------------------------------------------------------------------------------------
struct RunState {
    GpuTask& task;
    TaskGraph& g;
    Plan& plan;
    std::vector<DeviceRuntime*> slots;
    std::vector<int> slot_of_component;
    WorkerPool& pool;
    progress_callback_t progress_cb;
    void* user_data;
    std::any frontend_ctx;

    std::mutex mu;
    std::condition_variable cv;
    std::unordered_map<NodeIndex, std::any> avail;
    std::unordered_map<NodeIndex, void*> output_handles;
    std::vector<int> remaining;
    std::unordered_map<NodeIndex, int> refcnt;
    std::vector<std::vector<int>> ready_q;  // one queue per device slot; units without a device (host nodes, replicated loads) use queue 0
    bool aborted = false;                    // a dispatcher died: everybody leaves
    size_t done_units = 0;
    int in_flight = 0;  // units handed to the worker pool and not finished
    int feeding = 0;    // ... of which can still make GPU units ready (loads, host nodes; not the result drains)
    std::atomic<size_t> completed_nodes{0};
    std::atomic<bool> failed{false};
    std::string error;
    std::atomic<int64_t> last_progress_ns{0};
    std::atomic<long long> groups{0};

    // Result drains: a D2H copy is awaited by ONE helper thread (copies on the store stream finish in order), and only
    // then is the IMPORT work handed to the pool -- pool threads never block on the GPU (measured on cfg2 x1024: 15
    // threads spent 16 ms each of a 67 ms call inside cudaEventSynchronize).
    struct PendingDrain {
        cudaEvent_t ev;
        int device;
        int priority;
        std::function<void()> work;
    };
    std::deque<PendingDrain> drains;
    std::mutex drain_mu;
    std::condition_variable drain_cv;
    bool drain_stop = false;
    std::thread drainer;
    void drain_loop() {
        for (;;) {
            PendingDrain d;
            {
                std::unique_lock<std::mutex> lk(drain_mu);
                drain_cv.wait(lk, [&] { return drain_stop || !drains.empty(); });
                if (drains.empty()) return;
                d = std::move(drains.front());
                drains.pop_front();
            }
            cudaSetDevice(d.device);
            const long long t0 = now_ns();
            cudaEventSynchronize(d.ev);  // an error surfaces in the continuation's own CUDA calls
            task.ns_d2h_wait += now_ns() - t0;
            pool.submit(d.priority, std::move(d.work));
        }
    }
    void submit_drain(PendingDrain d) {
        {
            std::lock_guard<std::mutex> lk(drain_mu);
            if (!drainer.joinable()) drainer = std::thread([this] { drain_loop(); });
            drains.push_back(std::move(d));
        }
        drain_cv.notify_one();
    }
    void stop_drainer() {
        {
            std::lock_guard<std::mutex> lk(drain_mu);
            drain_stop = true;
        }
        drain_cv.notify_all();
        if (drainer.joinable()) drainer.join();
    }
    ~RunState() { stop_drainer(); }

    RunState(GpuTask& t, WorkerPool& wp, progress_callback_t cb, void* ud)
        : task(t), g(*t.graph), plan(t.plan), pool(wp), progress_cb(cb), user_data(ud) {}

    int slot_of(const Unit& u) const { return u.component >= 0 ? slot_of_component[u.component] : -1; }
    static bool feeds_gpu(const Unit& u) {
        return u.kind == UnitKind::LOAD || (u.kind == UnitKind::HOST && op_of(u.nodes[0]) != OperationType::IMPORT_FROM_ABI);
    }
    int queue_of(int unit) const { return std::max(0, slot_of(plan.units[unit])); }

    void fail(const std::string& what) {
        std::lock_guard<std::mutex> lk(mu);
        if (!failed.exchange(true)) error = what;
    }

    void report_progress(size_t done) {
        if (!progress_cb) return;
        const size_t total = plan.total_nodes;
        const int64_t now = std::chrono::steady_clock::now().time_since_epoch().count();
        const int64_t last = last_progress_ns.load(std::memory_order_relaxed);
        const int64_t interval = std::chrono::duration_cast<std::chrono::steady_clock::duration>(std::chrono::milliseconds(100)).count();
        if (done >= total || now - last >= interval) {
            last_progress_ns.store(now, std::memory_order_relaxed);
            progress_cb((int)done, (int)total, user_data);
        }
    }

    // publish a unit's output, release inputs nobody needs any more, wake successors
    void unit_done(const Unit& u, std::any output, bool from_pool) {
        std::vector<std::any> graveyard;  // destroyed outside the lock (device frees, C struct frees)
        {
            std::lock_guard<std::mutex> lk(mu);
            avail[u.output->index] = std::move(output);
            for (DatumNode* in : u.ext_inputs) {
                auto it = refcnt.find(in->index);
                if (it == refcnt.end()) continue;
                if (--it->second <= 0 && !in->is_input && !in->is_output) {
                    auto a = avail.find(in->index);
                    if (a != avail.end()) {
                        graveyard.push_back(std::move(a->second));
                        avail.erase(a);
                    }
                }
            }
            auto c = plan.consumers.find(u.output->index);
            if (c != plan.consumers.end())
                for (int v : c->second)
                    if (--remaining[v] == 0) ready_q[queue_of(v)].push_back(v);
            done_units++;
            if (from_pool) {
                in_flight--;
                if (feeds_gpu(u)) feeding--;
            }
        }
        cv.notify_all();
        const size_t done = completed_nodes.fetch_add(u.nodes.size()) + u.nodes.size();
        report_progress(done);
    }

    std::unordered_map<NodeIndex, std::any> gather_inputs(const Unit& u) {
        std::unordered_map<NodeIndex, std::any> in;
        std::lock_guard<std::mutex> lk(mu);
        for (DatumNode* d : u.ext_inputs) {
            auto it = avail.find(d->index);
            if (it == avail.end()) throw std::runtime_error("lsb runner: input '" + d->id + "' of node '" + u.nodes[0]->id + "' is not available");
            in[d->index] = it->second;
        }
        return in;
    }

    void run_host_unit(const Unit& u);
    void run_load_unit(const Unit& u);
    void start_store_unit(const Unit& u);
    void flush_gpu(std::vector<int>& pending);
    void exec_group(int slot, const std::vector<int>& members, const struct GpuCall& rep, std::vector<struct GpuCall>& calls);
    struct GpuCall describe(const Unit& u, const std::unordered_map<NodeIndex, std::any>& in, int slot);
    void dispatch(int q);
};

// worker-pool priorities: results drain first, then input staging in consumer order, then other host nodes
static const int kDrainPriority = 2000000;
static const int kLoadPriority = 1000000;

// ---- host nodes: EXPORT_TO_ABI, IMPORT_FROM_ABI, custom (cpu_task_utils.h:420-506) ----
void RunState::run_host_unit(const Unit& u) {
    const ComputeNode& node = *u.nodes[0];
    std::any output;
    try {
        if (failed.load()) throw std::runtime_error("cancelled");
        if (!node.executor) throw std::runtime_error("lsb runner: no executor bound for node '" + node.id + "'");
        auto in = gather_inputs(u);
        ExecutionContext ctx;
        ctx.context = frontend_ctx;
        if (op_of(&node) == OperationType::IMPORT_FROM_ABI) {
            auto it = output_handles.find(u.output->index);
            if (it != output_handles.end()) ctx.other_args.push_back(std::any(it->second));
        }
        node.executor(ctx, in, output, node);
    } catch (const std::exception& e) {
        if (std::string(e.what()) != "cancelled") fail("node '" + node.id + "': " + e.what());
    } catch (...) {
        fail("node '" + node.id + "': unknown exception");
    }
    unit_done(u, std::move(output), true);
}

// ---- LOAD_TO_BACKEND (gpu_abi_bridge_executors.h:213-310): C struct -> contiguous device object ----
namespace {
struct LimbList {
    std::vector<const uint64_t*> limbs;  // in device order
    int n = 0, level = 0, npoly = 0, limbs_per_poly = 0;
};
void collect_poly(const CPolynomial& p, LimbList& out) {
    for (int i = 0; i < p.n_component; i++) {
        if (out.n == 0) out.n = p.components[i].n;
        LSB_CHECK(p.components[i].n == out.n && p.components[i].data, "lsb runner: malformed CComponent");
        out.limbs.push_back(p.components[i].data);
    }
}
void collect_ksk(const CKeySwitchKey& k, LimbList& out) {
    LSB_CHECK(k.n_public_key >= 1 && k.public_keys, "lsb runner: empty key-switch key");
    out.level = k.public_keys[0].level;
    out.npoly = 0;
    for (int d = 0; d < k.n_public_key; d++) {
        const CPublicKey& pk = k.public_keys[d];
        LSB_CHECK(pk.degree == 1 && pk.level == out.level, "lsb runner: malformed public key in key-switch key");
        for (int j = 0; j < 2; j++) {
            out.limbs_per_poly = pk.polys[j].n_component;
            collect_poly(pk.polys[j], out);
        }
    }
}
}  // namespace

// words of the object a LOAD unit is about to receive, from the datum's metadata (an estimate: the arena falls back to
// malloc when the executor allocates more, and the staging copy takes over)
static size_t expected_words(const TaskGraph& g, const DatumNode* d) {
    if (!d->fhe_prop.has_value()) return 0;
    const size_t n = (size_t)g.param.n, l = (size_t)d->fhe_prop->level + 1, k = g.param.p.size();
    switch (d->datum_type) {
        case TYPE_CIPHERTEXT: return (size_t)(d->fhe_prop->degree + 1) * l * n;
        case TYPE_PLAINTEXT: return (d->fhe_prop->p.has_value() && d->fhe_prop->p->is_ringt) ? n : l * n;
        case TYPE_RELIN_KEY:
        case TYPE_GALOIS_KEY:
        case TYPE_SWITCH_KEY: return k ? ((l + k - 1) / k) * 2 * (l + k) * n : 0;
        default: return 0;
    }
}

void RunState::run_load_unit(const Unit& u) {
    const ComputeNode& node = *u.nodes.back();
    std::any output;
    try {
        if (failed.load()) throw std::runtime_error("cancelled");
        const DatumNode* src = node.input_nodes[0];
        LSB_CHECK(src->fhe_prop.has_value(), "Input node missing FHE properties");
        auto in = gather_inputs(u);
        // ---- device-resident key cache: a key handle seen before on this task is not exported or uploaded again ----
        const bool cacheable = task.key_cache_on && is_key_type(src->datum_type) && u.ext_inputs.size() == 1 && u.ext_inputs[0]->is_input;
        std::pair<const void*, NodeIndex> cache_key{nullptr, u.output->index};
        if (cacheable) {
            if (auto* hp = std::any_cast<std::shared_ptr<void>>(&in.at(u.ext_inputs[0]->index))) cache_key.first = hp->get();
            if (cache_key.first) {
                std::lock_guard<std::mutex> lk(task.key_mu);
                auto it = task.key_cache.find(cache_key);
                if (it != task.key_cache.end()) {
                    output = it->second;
                    task.stat_key_cached += (long long)expected_words(g, src) * 8;
                }
            }
            if (output.has_value()) {
                unit_done(u, std::move(output), true);
                return;
            }
        }
        // ---- which device slots want this object ----
        const int my_slot = slot_of(u);
        std::vector<char> wanted(slots.size(), my_slot < 0 ? 0 : 1);
        if (my_slot < 0) {  // a replicated object (key, plaintext) goes to every device that runs one of its consumers
            auto cons = plan.consumers.find(u.output->index);
            if (cons != plan.consumers.end())
                for (int v : cons->second) {
                    const int cs = slot_of(plan.units[v]);
                    if (cs >= 0) wanted[cs] = 1;
                    else std::fill(wanted.begin(), wanted.end(), 1);
                }
        }
        int owner = my_slot;
        for (int s2 = 0; owner < 0 && s2 < (int)slots.size(); s2++)
            if (wanted[s2]) owner = s2;
        // ---- EXPORT_TO_ABI fused into this unit: run the bound executor here, lending it pinned staging memory ----
        std::any exported;
        lsr::ExportArena arena;
        uint64_t arena_ticket = 0;
        const size_t want_words = expected_words(g, src);
        static const bool zero_copy = [] {
            const char* e = getenv("LSB_EXPORT_ZERO_COPY");
            return !(e && atoi(e) == 0);
        }();
        if (u.nodes.size() == 2) {
            const ComputeNode& ex = *u.nodes[0];
            if (!ex.executor) throw std::runtime_error("lsb runner: no executor bound for node '" + ex.id + "'");
            ExecutionContext ctx;
            ctx.context = frontend_ctx;
            const bool lend = zero_copy && owner >= 0 && want_words > 0 && want_words * 8 <= slots[owner]->up.cap / 2;
            if (lend) {
                const long long t0 = now_ns();
                arena.base = slots[owner]->up.acquire(want_words * 8, &arena_ticket);
                task.ns_ring_wait += now_ns() - t0;
                arena.cap = want_words * 8;
                lsr::export_arena_begin(&arena);
            }
            const long long t_ex = now_ns();
            try {
                ex.executor(ctx, in, exported, ex);
                task.ns_export += now_ns() - t_ex;
            } catch (...) {
                if (lend) {
                    lsr::export_arena_end();
                    slots[owner]->up.release(arena_ticket);
                }
                throw;
            }
            if (lend) lsr::export_arena_end();
        }
        const std::any& cs = u.nodes.size() == 2 ? exported : in.at(src->index);
        LimbList ll;
        std::shared_ptr<void> keepalive;
        switch (src->datum_type) {
            case TYPE_CIPHERTEXT: {
                auto c = std::any_cast<std::shared_ptr<CCiphertext>>(cs);
                ll.level = c->level;
                ll.npoly = c->degree + 1;
                for (int p = 0; p <= c->degree; p++) {
                    ll.limbs_per_poly = c->polys[p].n_component;
                    collect_poly(c->polys[p], ll);
                }
                keepalive = c;
                break;
            }
            case TYPE_PLAINTEXT: {
                auto c = std::any_cast<std::shared_ptr<CPlaintext>>(cs);
                ll.level = c->level;
                ll.npoly = 1;
                ll.limbs_per_poly = c->poly.n_component;
                collect_poly(c->poly, ll);
                keepalive = c;
                break;
            }
            case TYPE_RELIN_KEY: {
                auto c = std::any_cast<std::shared_ptr<CRelinKey>>(cs);
                collect_ksk(*c, ll);
                keepalive = c;
                break;
            }
            case TYPE_SWITCH_KEY: {
                auto c = std::any_cast<std::shared_ptr<CKeySwitchKey>>(cs);
                collect_ksk(*c, ll);
                keepalive = c;
                break;
            }
            case TYPE_GALOIS_KEY: {
                auto c = std::any_cast<std::shared_ptr<CGaloisKey>>(cs);
                const uint64_t want = src->fhe_prop->p.has_value() ? src->fhe_prop->p->galois_element : 0;
                int found = -1;
                for (int i = 0; i < c->n_key_switch_key; i++)
                    if (c->galois_elements[i] == want) found = i;
                if (found < 0 && c->n_key_switch_key == 1 && want == 0) found = 0;
                LSB_CHECK(found >= 0, "lsb runner: Galois key for element " + std::to_string(want) + " missing from the exported CGaloisKey");
                collect_ksk(c->key_switch_keys[found], ll);
                keepalive = c;
                break;
            }
            default: throw std::runtime_error("Unsupported data type in H2D transfer");
        }
        LSB_CHECK(ll.n == g.param.n, "lsb runner: ring degree of '" + src->id + "' does not match the task parameter");
        const size_t words = ll.limbs.size() * (size_t)ll.n;
        // did the executor fill the lent region limb after limb, in the order the device layout wants?
        bool in_place = arena.base && !arena.overflow && arena.used == words * 8;
        for (size_t i = 0; in_place && i < ll.limbs.size(); i++)
            in_place = (const uint8_t*)ll.limbs[i] == arena.base + i * (size_t)ll.n * 8;
        DatumPtr datum = std::make_shared<DeviceDatum>();
        datum->on.resize(slots.size());
        for (int s = 0; s < (int)slots.size(); s++) {
            if (!wanted[s] || (my_slot >= 0 && s != my_slot)) continue;
            DeviceRuntime* rt = slots[s];
            LSB_CUDA(cudaSetDevice(rt->device));
            uint64_t ticket = 0;
            uint8_t* stage = nullptr;
            if (s == owner && arena.base && (in_place || (arena.used == 0 && words * 8 <= arena.cap))) {
                // the lent region IS the staging buffer: already filled (zero copy), or untouched and reused for the copy
                stage = arena.base;
                ticket = arena_ticket;
                arena.base = nullptr;
                if (in_place) {
                    task.stat_zero_copy += (long long)words * 8;
                } else {
                    const long long t0 = now_ns();
                    for (size_t i = 0; i < ll.limbs.size(); i++) lsr_stream_copy(stage + i * (size_t)ll.n * 8, ll.limbs[i], (size_t)ll.n * 8);
                    task.ns_stage += now_ns() - t0;
                }
            } else {
                long long t0 = now_ns();
                stage = rt->up.acquire(words * 8, &ticket);
                task.ns_ring_wait += now_ns() - t0;
                t0 = now_ns();
                for (size_t i = 0; i < ll.limbs.size(); i++) lsr_stream_copy(stage + i * (size_t)ll.n * 8, ll.limbs[i], (size_t)ll.n * 8);
                task.ns_stage += now_ns() - t0;
            }
            cudaEvent_t ring_ev = rt->up.new_event();
            DevObjPtr obj;
            {
                std::lock_guard<std::mutex> lk(rt->mu);
                cudaStream_t up = rt->next_h2d();
                obj = alloc_device_object(rt, words, up);
                LSB_CUDA(cudaMemcpyAsync(obj->ptr, stage, words * 8, cudaMemcpyHostToDevice, up));
                LSB_CUDA(cudaEventRecord(ring_ev, up));
                LSB_CUDA(cudaEventCreateWithFlags(&obj->ready, cudaEventDisableTiming));
                LSB_CUDA(cudaEventRecord(obj->ready, up));
            }
            rt->up.commit(ticket, ring_ev);
            task.stat_h2d += (long long)words * 8;
            obj->level = ll.level;
            obj->npoly = ll.npoly;
            obj->limbs = ll.limbs_per_poly;
            datum->on[s] = obj;
        }
        if (arena.base) {
            // lent region not used as staging (mismatching layout): its limbs stay valid until the exported struct dies
            // with `keepalive` at the end of this scope, then the region goes back to the ring
            keepalive.reset();
            exported.reset();
            slots[owner]->up.release(arena_ticket);
        }
        output = datum;
        if (cacheable && cache_key.first) {
            std::lock_guard<std::mutex> lk(task.key_mu);
            task.key_cache[cache_key] = datum;
        }
    } catch (const std::exception& e) {
        if (std::string(e.what()) != "cancelled") fail("node '" + node.id + "': " + e.what());
    }
    unit_done(u, std::move(output), true);
}

// ---- STORE_FROM_BACKEND (gpu_abi_bridge_executors.h:326-363): device object -> freshly allocated CCiphertext ----
void RunState::start_store_unit(const Unit& u) {
    const ComputeNode& node = *u.nodes[0];
    try {
        if (failed.load()) throw std::runtime_error("cancelled");
        const DatumNode* src = node.input_nodes[0];
        LSB_CHECK(src->fhe_prop.has_value(), "Input node missing FHE properties");
        if (src->datum_type != TYPE_CIPHERTEXT) throw std::runtime_error("Unsupported data type for D2H transfer");
        auto in = gather_inputs(u);
        DatumPtr datum = std::any_cast<DatumPtr>(in.at(src->index));
        const int s = slot_of(u);
        DevObjPtr obj = datum->on.at(s);
        LSB_CHECK(obj, "lsb runner: output '" + src->id + "' is not resident on its device");
        DeviceRuntime* rt = slots[s];
        LSB_CUDA(cudaSetDevice(rt->device));
        uint64_t ticket = 0;
        uint8_t* stage = rt->down.acquire(obj->words * 8, &ticket);
        cudaEvent_t done_ev;
        LSB_CUDA(cudaEventCreateWithFlags(&done_ev, cudaEventDisableTiming));
        {
            std::lock_guard<std::mutex> lk(rt->mu);
            // everything that produced obj was enqueued on the compute stream by this (dispatcher) thread
            LSB_CUDA(cudaEventRecord(done_ev, rt->compute));
            LSB_CUDA(cudaStreamWaitEvent(rt->d2h, done_ev, 0));
            if (obj->ready) LSB_CUDA(cudaStreamWaitEvent(rt->d2h, obj->ready, 0));
            LSB_CUDA(cudaMemcpyAsync(stage, obj->ptr, obj->words * 8, cudaMemcpyDeviceToHost, rt->d2h));
            task.stat_d2h += (long long)obj->words * 8;
            LSB_CUDA(cudaEventRecord(done_ev, rt->d2h));
        }
        {
            std::lock_guard<std::mutex> lk(mu);
            in_flight++;
        }
        const Unit* up = &u;
        // results drain ahead of further input staging
        PendingDrain pd;
        pd.ev = done_ev;
        pd.device = rt->device;
        pd.priority = kDrainPriority + u.priority;
        pd.work = [this, up, obj, rt, stage, ticket, done_ev] {
            std::any output;
            const int n = g.param.n;
            try {
                LSB_CUDA(cudaSetDevice(rt->device));
                LSB_CUDA(cudaEventSynchronize(done_ev));  // already complete (awaited by the drain thread)
                const long long t_imp = now_ns();
                struct ImportTimer {
                    GpuTask& t;
                    long long t0;
                    ~ImportTimer() { t.ns_import += now_ns() - t0; }
                } import_timer{task, t_imp};
                if (up->nodes.size() == 2) {
                    // fused IMPORT_FROM_ABI: hand the bound executor a CCiphertext whose limbs ARE the pinned staging
                    // region (valid for the duration of the call; the struct is owned by the runner, wrapper.h ownership
                    // rules are unchanged for the caller's handles)
                    const ComputeNode& imp = *up->nodes[1];
                    if (!imp.executor) throw std::runtime_error("lsb runner: no executor bound for node '" + imp.id + "'");
                    CCiphertext view;
                    view.level = obj->level;
                    view.degree = obj->npoly - 1;
                    std::vector<CPolynomial> polys(obj->npoly);
                    std::vector<CComponent> comps((size_t)obj->npoly * (obj->level + 1));
                    size_t k = 0;
                    for (int pi = 0; pi < obj->npoly; pi++) {
                        polys[pi].n_component = obj->level + 1;
                        polys[pi].components = &comps[k];
                        for (int li = 0; li <= obj->level; li++, k++) {
                            comps[k].n = n;
                            comps[k].data = reinterpret_cast<uint64_t*>(stage + k * (size_t)n * 8);
                        }
                    }
                    view.polys = polys.data();
                    std::unordered_map<NodeIndex, std::any> in;
                    in[imp.input_nodes[0]->index] = std::shared_ptr<CCiphertext>(&view, [](CCiphertext*) {});
                    ExecutionContext ctx;
                    ctx.context = frontend_ctx;
                    auto it = output_handles.find(up->output->index);
                    if (it != output_handles.end()) ctx.other_args.push_back(std::any(it->second));
                    imp.executor(ctx, in, output, imp);
                } else {
                    CCiphertext* c = (CCiphertext*)malloc(sizeof(CCiphertext));
                    alloc_ciphertext(c, obj->npoly - 1, obj->level, n);
                    std::shared_ptr<CCiphertext> sp(c, [](CCiphertext* p) {
                        free_ciphertext(p);
                        free(p);
                    });
                    size_t k = 0;
                    for (int pi = 0; pi < obj->npoly; pi++)
                        for (int li = 0; li <= obj->level; li++, k++)
                            std::memcpy(c->polys[pi].components[li].data, stage + k * (size_t)n * 8, (size_t)n * 8);
                    output = sp;
                }
            } catch (const std::exception& e) {
                fail("node '" + up->nodes.back()->id + "': " + e.what());
            } catch (...) {
                fail("node '" + up->nodes.back()->id + "': unknown exception");
            }
            rt->down.release(ticket);  // the staging region is free again
            cudaEventDestroy(done_ev);
            unit_done(*up, std::move(output), true);
        };
        submit_drain(std::move(pd));
        return;
    } catch (const std::exception& e) {
        if (std::string(e.what()) != "cancelled") fail("node '" + node.id + "': " + e.what());
    }
    unit_done(u, std::any(), false);
}


// ---------------------------------------------------------------------------------------------
// GPU nodes: describe -> group by signature -> one batched launch train per group
// (replaces the per-node HEonGPU calls of mega_ag_executors_gpu.cu:71-426)
// ---------------------------------------------------------------------------------------------
enum CallKind {
    K_ADD_CC, K_SUB_CC, K_ADD_CP, K_ADD_CPR, K_SUB_PC, K_SUB_PRC, K_NEG, K_MULT_CC, K_MUL_CP, K_MUL_CPR, K_RELIN, K_RESCALE,
    K_DROP, K_ROTATE, K_MAC, K_MULT_RELIN, K_MULT_RELIN_RESCALE, K_SUM, K_BOOTSTRAP
};
struct GpuCall {
    int kind = 0, level = 0, npoly = 2, aux = 0, aux2 = 0;  // aux: subtract flag / ringt flag / sum_cnt ; aux2: has partial sum
    const u64* key = nullptr;
    int key_level = 0;
    uint64_t gal = 0;
    std::vector<DevObjPtr> operands;  // inputs in the canonical order of the kind
    std::vector<DevObjPtr> keepalive;
    std::shared_ptr<lsb::BootstrapKeys> btp_keys;  // K_BOOTSTRAP: rlk, Galois keys by element, swk_dts, swk_std
    int out_level = 0, out_npoly = 2;
    std::string signature(int slot) const {
        char buf[160];
        snprintf(buf, sizeof buf, "%d|%d|%d|%d|%d|%d|%p|%d|%llu|%zu|%d|%d", slot, kind, level, npoly, aux, aux2, (const void*)key,
                 key_level, (unsigned long long)gal, operands.size(), out_level, out_npoly);
        return buf;
    }
};

// bootstrapping parameters as the frontend writes them into mega_ag.json (custom_task.py CkksBtpParam.to_dict; the same
// fields init_gpu_context reads, gpu_wrapper.cu:92-116)
static lsb::btp::Params btp_params(const TaskGraph& g) {
    const Json& j = g.parameter;
    auto geti = [&](const char* k) {
        LSB_CHECK(j.contains(k), std::string("mega_ag: parameter '") + k + "' is missing");
        return j.at(k).as_int();
    };
    auto getd = [&](const char* k, double dflt) { return j.contains(k) ? j.at(k).as_double() : dflt; };
    lsb::btp::Params P;
    P.logN = 0;
    while ((1 << P.logN) < g.param.n) P.logN++;
    const int slots = j.contains("slots") ? j.at("slots").as_int() : g.param.n / 2;
    P.log_slots = 0;
    while ((1 << P.log_slots) < slots) P.log_slots++;
    P.q.assign(g.param.q.begin(), g.param.q.begin() + g.param.max_level + 1);
    P.p = g.param.p;
    P.scale = getd("scale", 0.0);
    P.cts_start = geti("btp_cts_start_level");
    P.cts_depth = geti("btp_cts_depth");
    P.cts_ratio = getd("btp_cts_bsgs_ratio", 2.0);
    P.evalmod_start = geti("btp_eval_mod_start_level");
    P.message_ratio = getd("btp_eval_mod_message_ratio", 256.0);
    P.K = geti("btp_eval_mod_k");
    P.sine_deg = geti("btp_eval_mod_sine_deg");
    P.double_angle = geti("btp_eval_mod_double_angle");
    LSB_CHECK(!j.contains("btp_eval_mod_arcsine_deg") || j.at("btp_eval_mod_arcsine_deg").as_int() == 0,
              "lsb runner: bootstrapping with an arcsine correction is not supported");
    LSB_CHECK(!j.contains("btp_eval_mod_sine_type") || j.at("btp_eval_mod_sine_type").as_string() == "Cos1",
              "lsb runner: only the Cos1 EvalMod variant is supported");
    P.stc_start = geti("btp_stc_start_level");
    P.stc_depth = geti("btp_stc_depth");
    P.stc_ratio = getd("btp_stc_bsgs_ratio", 2.0);
    P.out_level = geti("btp_output_level");
    return P;
}

namespace {
enum PtKind { PT_NONE, PT_PLAIN, PT_RINGT };
PtKind pt_kind(const DatumNode* d) {
    if (d->datum_type != TYPE_PLAINTEXT) return PT_NONE;
    return (d->fhe_prop->p.has_value() && d->fhe_prop->p->is_ringt) ? PT_RINGT : PT_PLAIN;
}
uint64_t galois_for_step(int step, int n) {
    // frontend/custom_task.py:139-141: 5^step mod 2N, step taken modulo the column count N/2
    const uint64_t m = 2ull * (uint64_t)n;
    long long e = step % (n / 2);
    if (e < 0) e += n / 2;
    uint64_t r = 1, b = 5 % m;
    while (e) {
        if (e & 1) r = r * b % m;
        b = b * b % m;
        e >>= 1;
    }
    return r;
}
}  // namespace

GpuCall RunState::describe(const Unit& u, const std::unordered_map<NodeIndex, std::any>& in, int slot) {
    GpuCall c;
    auto dev = [&](const DatumNode* d) -> DevObjPtr {
        DatumPtr p = std::any_cast<DatumPtr>(in.at(d->index));
        LSB_CHECK(p && (int)p->on.size() > slot && p->on[slot], "lsb runner: '" + d->id + "' is not resident on device slot " + std::to_string(slot));
        return p->on[slot];
    };
    const ComputeNode& n0 = *u.nodes[0];
    const auto& ins = n0.input_nodes;
    auto need = [&](size_t k) { LSB_CHECK(ins.size() >= k, "mega_ag: node '" + n0.id + "' has too few inputs"); };
    const DatumNode* out = u.output;
    LSB_CHECK(out->fhe_prop.has_value(), "mega_ag: output of '" + n0.id + "' has no FHE properties");
    c.out_level = out->fhe_prop->level;
    c.out_npoly = out->fhe_prop->degree + 1;
    if (!u.sum_terms.empty()) {  // re-associated sum: operands = ciphertexts (product terms first), then their plaintexts
        c.kind = K_SUM;
        c.level = c.out_level;
        c.npoly = c.out_npoly;
        for (auto& t : u.sum_terms) c.aux += t.second ? 1 : 0;
        for (auto& t : u.sum_terms) c.operands.push_back(dev(t.first));
        for (auto& t : u.sum_terms)
            if (t.second) c.operands.push_back(dev(t.second));
        return c;
    }
    if (u.nodes.size() > 1) {  // fused mult -> relin [-> rescale]
        need(1);
        const ComputeNode& relin = *u.nodes[1];
        c.kind = u.nodes.size() == 3 ? K_MULT_RELIN_RESCALE : K_MULT_RELIN;
        c.level = ins[0]->fhe_prop->level;
        c.operands = {dev(ins[0]), dev(ins.size() > 1 ? ins[1] : ins[0])};
        DevObjPtr k = dev(relin.input_nodes.at(1));
        c.key = k->ptr;
        c.key_level = k->level;
        c.keepalive.push_back(k);
        return c;
    }
    const OperationType op = op_of(&n0);
    switch (op) {
        case OperationType::ADD:
        case OperationType::SUB: {
            need(1);
            const bool sub = op == OperationType::SUB;
            if (ins.size() == 1) {  // x (op) x: the frontend emits ONE edge (mega_ag_executors_gpu.cu:72-80, 124-132)
                LSB_CHECK(pt_kind(ins[0]) == PT_NONE, "lsb runner: add/sub of a single plaintext is not a ciphertext operation");
                c.kind = sub ? K_SUB_CC : K_ADD_CC;
                c.level = ins[0]->fhe_prop->level;
                c.npoly = ins[0]->fhe_prop->degree + 1;
                c.operands = {dev(ins[0]), dev(ins[0])};
                break;
            }
            const PtKind p0 = pt_kind(ins[0]), p1 = pt_kind(ins[1]);
            if (p0 == PT_NONE && p1 == PT_NONE) {
                c.kind = sub ? K_SUB_CC : K_ADD_CC;
                c.level = ins[0]->fhe_prop->level;
                c.npoly = ins[0]->fhe_prop->degree + 1;
                LSB_CHECK(ins[1]->fhe_prop->degree + 1 == c.npoly && ins[1]->fhe_prop->level == c.level,
                          "lsb runner: add/sub operands of '" + n0.id + "' differ in level or degree");
                c.operands = {dev(ins[0]), dev(ins[1])};
            } else if (p0 == PT_NONE) {
                c.kind = p1 == PT_RINGT ? K_ADD_CPR : K_ADD_CP;
                c.aux = sub ? 1 : 0;
                c.level = ins[0]->fhe_prop->level;
                c.npoly = ins[0]->fhe_prop->degree + 1;
                c.operands = {dev(ins[0]), dev(ins[1])};
            } else {
                LSB_CHECK(p1 == PT_NONE, "lsb runner: plaintext (op) plaintext is not a ciphertext operation");
                c.kind = !sub ? (p0 == PT_RINGT ? K_ADD_CPR : K_ADD_CP) : (p0 == PT_RINGT ? K_SUB_PRC : K_SUB_PC);
                c.level = ins[1]->fhe_prop->level;
                c.npoly = ins[1]->fhe_prop->degree + 1;
                c.operands = {dev(ins[1]), dev(ins[0])};
            }
            break;
        }
        case OperationType::NEGATE:
            need(1);
            c.kind = K_NEG;
            c.level = ins[0]->fhe_prop->level;
            c.npoly = ins[0]->fhe_prop->degree + 1;
            c.operands = {dev(ins[0])};
            break;
        case OperationType::MULTIPLY: {
            need(1);
            if (ins.size() == 1) {  // square
                LSB_CHECK(pt_kind(ins[0]) == PT_NONE && ins[0]->fhe_prop->degree == 1, "lsb runner: square needs a degree-1 ciphertext");
                c.kind = K_MULT_CC;
                c.level = ins[0]->fhe_prop->level;
                c.operands = {dev(ins[0]), dev(ins[0])};
                break;
            }
            const PtKind p0 = pt_kind(ins[0]), p1 = pt_kind(ins[1]);
            if (p0 == PT_NONE && p1 == PT_NONE) {
                LSB_CHECK(ins[0]->fhe_prop->degree == 1 && ins[1]->fhe_prop->degree == 1, "lsb runner: ct x ct needs degree-1 operands");
                c.kind = K_MULT_CC;
                c.level = ins[0]->fhe_prop->level;
                c.operands = {dev(ins[0]), dev(ins[1])};
            } else {
                const int ci = p0 == PT_NONE ? 0 : 1, pi = 1 - ci;
                LSB_CHECK(pt_kind(ins[ci]) == PT_NONE, "lsb runner: plaintext x plaintext is not a ciphertext operation");
                const PtKind pk = pt_kind(ins[pi]);
                if (pk == PT_PLAIN && g.algo == ALGO_BFV)
                    LSB_CHECK(ins[pi]->fhe_prop->is_ntt, "lsb runner: BFV ct x pt needs a pt_mul operand (ct x scaled plaintext is CPU-only in the reference, "
                                                          "unittests/test_cpu_bfv.cpp:304)");
                c.kind = pk == PT_RINGT ? K_MUL_CPR : K_MUL_CP;
                c.level = ins[ci]->fhe_prop->level;
                c.npoly = ins[ci]->fhe_prop->degree + 1;
                c.operands = {dev(ins[ci]), dev(ins[pi])};
            }
            break;
        }
        case OperationType::RELINEARIZE: {
            need(2);
            c.kind = K_RELIN;
            c.level = ins[0]->fhe_prop->level;
            c.operands = {dev(ins[0])};
            DevObjPtr k = dev(ins[1]);
            c.key = k->ptr;
            c.key_level = k->level;
            c.keepalive.push_back(k);
            break;
        }
        case OperationType::RESCALE:
        case OperationType::DROP_LEVEL:
            need(1);
            c.kind = op == OperationType::RESCALE ? K_RESCALE : K_DROP;
            c.level = ins[0]->fhe_prop->level;
            c.npoly = ins[0]->fhe_prop->degree + 1;
            c.operands = {dev(ins[0])};
            break;
        case OperationType::ROTATE_COL:
        case OperationType::ROTATE_ROW: {
            need(2);
            c.kind = K_ROTATE;
            c.level = ins[0]->fhe_prop->level;
            c.operands = {dev(ins[0])};
            DevObjPtr k = dev(ins[1]);
            c.key = k->ptr;
            c.key_level = k->level;
            c.keepalive.push_back(k);
            // prefer the key datum's element (SEAL-flavoured graphs use generator 3)
            uint64_t gal = ins[1]->fhe_prop->p.has_value() ? ins[1]->fhe_prop->p->galois_element : 0;
            if (!gal)
                gal = op == OperationType::ROTATE_ROW ? 2ull * g.param.n - 1
                                                     : galois_for_step(n0.fhe_prop->p.has_value() ? n0.fhe_prop->p->rotation_step : 0, g.param.n);
            c.gal = gal;
            break;
        }
        case OperationType::MAC_WO_PARTIAL_SUM:
        case OperationType::MAC_W_PARTIAL_SUM: {
            // inputs [ct_0..ct_{n-1}, (partial sum), pt_0..pt_{n-1}] (mega_ag_executors_gpu.cu:300,359)
            const bool with_acc = op == OperationType::MAC_W_PARTIAL_SUM;
            const int cnt = n0.fhe_prop->p.has_value() ? n0.fhe_prop->p->sum_cnt : 0;
            LSB_CHECK(cnt >= 1 && (int)ins.size() == 2 * cnt + (with_acc ? 1 : 0), "mega_ag: MAC node '" + n0.id + "' has an inconsistent input list");
            c.kind = K_MAC;
            c.level = ins[0]->fhe_prop->level;
            c.aux2 = with_acc ? 1 : 0;
            const DatumNode* pt0 = ins[cnt + (with_acc ? 1 : 0)];
            c.aux = pt_kind(pt0) == PT_RINGT ? 1 : 0;
            for (int i = 0; i < cnt; i++) {
                const DatumNode* pti = ins[cnt + (with_acc ? 1 : 0) + i];
                LSB_CHECK(pt_kind(pti) == pt_kind(pt0) && pt_kind(pti) != PT_NONE, "mega_ag: MAC node '" + n0.id + "' mixes plaintext kinds");
                // reference: 'Multiply with plaintext only supported for CKKS' (mega_ag_executors_gpu.cu:202-213); here a BFV
                // plaintext is accepted when it is a pt_mul (NTT domain), never a coefficient-domain scaled plaintext
                if (g.algo == ALGO_BFV && pt_kind(pti) == PT_PLAIN)
                    LSB_CHECK(pti->fhe_prop->is_ntt, "lsb runner: BFV multiply-accumulate needs pt_ringt or pt_mul operands (node '" + n0.id + "')");
            }
            c.npoly = cnt;  // number of terms
            for (int i = 0; i < (int)ins.size(); i++) c.operands.push_back(dev(ins[i]));
            break;
        }
        case OperationType::BOOTSTRAP: {
            // inputs [ct, rlk, glk..., swk_dts, swk_std] (mega_ag_executors_gpu.cu:415-419, custom_task.py:1961-1996)
            need(5);
            LSB_CHECK(g.algo == ALGO_CKKS && g.param.has_btp, "lsb runner: bootstrap needs a CKKS task with btp_* parameters (node '" + n0.id + "')");
            LSB_CHECK(ins[0]->fhe_prop->level == 0 && ins[0]->fhe_prop->degree == 1, "lsb runner: bootstrap input must be a level-0 ciphertext");
            c.kind = K_BOOTSTRAP;
            c.level = 0;
            c.operands = {dev(ins[0])};
            auto keys = std::make_shared<lsb::BootstrapKeys>();
            int n_swk = 0;
            for (size_t i = 1; i < ins.size(); i++) {
                DevObjPtr k = dev(ins[i]);
                c.keepalive.push_back(k);
                switch (ins[i]->datum_type) {
                    case TYPE_RELIN_KEY:
                        keys->rlk = k->ptr;
                        keys->key_level = k->level;
                        break;
                    case TYPE_GALOIS_KEY:
                        LSB_CHECK(ins[i]->fhe_prop->p.has_value() && ins[i]->fhe_prop->p->galois_element, "mega_ag: Galois key '" + ins[i]->id + "' carries no element");
                        keys->gal.push_back(ins[i]->fhe_prop->p->galois_element);
                        keys->glk.push_back(k->ptr);
                        LSB_CHECK(k->level == g.param.max_level, "lsb runner: bootstrap Galois keys must be at the top level");
                        break;
                    case TYPE_SWITCH_KEY:
                        // the last two inputs: swk_dts (level 0) then swk_std (top level)
                        if (n_swk == 0) {
                            keys->swk_dts = k->ptr;
                            keys->swk_dts_level = k->level;
                        } else {
                            keys->swk_std = k->ptr;
                            LSB_CHECK(k->level == g.param.max_level, "lsb runner: swk_std must be at the top level");
                        }
                        n_swk++;
                        break;
                    default: throw std::runtime_error("mega_ag: unexpected input '" + ins[i]->id + "' of bootstrap node '" + n0.id + "'");
                }
            }
            LSB_CHECK(keys->rlk && keys->swk_dts && keys->swk_std && n_swk == 2 && !keys->glk.empty(), "mega_ag: bootstrap node '" + n0.id + "' lacks a key input");
            LSB_CHECK(keys->key_level == g.param.max_level, "lsb runner: the bootstrap relinearisation key must be at the top level");
            c.key = keys->rlk;
            c.key_level = keys->key_level;
            c.btp_keys = keys;
            break;
        }
        default: throw std::runtime_error("lsb runner: unsupported operation in node '" + n0.id + "'");
    }
    return c;
}

// device pointer table for one operand of a batched launch
static u64** push_table(DeviceRuntime* rt, const std::vector<u64*>& ptrs) {
    const size_t n = ptrs.size();
    LSB_CHECK(n <= rt->tab_cap, "lsb runner: batch too large for the pointer-table arena");
    if (rt->tab_head + n > rt->tab_cap) {
        // wrap: earlier tables may still be read by queued copies/kernels
        LSB_CUDA(cudaStreamSynchronize(rt->compute));
        rt->tab_head = 0;
    }
    u64** h = rt->tab_host + rt->tab_head;
    u64** d = rt->tab_dev + rt->tab_head;
    for (size_t i = 0; i < n; i++) h[i] = ptrs[i];
    LSB_CUDA(cudaMemcpyAsync(d, h, n * sizeof(u64*), cudaMemcpyHostToDevice, rt->compute));
    rt->tab_head += n;
    return d;
}

void RunState::exec_group(int slot, const std::vector<int>& members, const GpuCall& rep, std::vector<GpuCall>& calls) {
    DeviceRuntime* rt = slots[slot];
    lsb::Context& ctx = *rt->ctx;
    LSB_CUDA(cudaSetDevice(rt->device));
    const int nb = (int)members.size();
    const size_t N = ctx.N;
    // inputs produced on the copy stream: make the compute stream wait for them once
    for (GpuCall& c : calls) {
        for (auto& o : c.operands)
            if (o->ready) LSB_CUDA(cudaStreamWaitEvent(rt->compute, o->ready, 0));
        for (auto& o : c.keepalive)
            if (o->ready) LSB_CUDA(cudaStreamWaitEvent(rt->compute, o->ready, 0));
    }
    LSB_CHECK(rep.out_level >= 0 && rep.out_level < ctx.nq, "lsb runner: output level outside the modulus chain");
    LSB_CHECK(rep.level >= 0 && rep.level < ctx.nq, "lsb runner: ciphertext level outside the modulus chain");
    const size_t out_words = (size_t)rep.out_npoly * (rep.out_level + 1) * N;
    std::vector<DevObjPtr> outs(nb);
    std::vector<u64*> optr(nb);
    for (int i = 0; i < nb; i++) {
        outs[i] = alloc_device_object(rt, out_words, rt->compute);
        outs[i]->level = rep.out_level;
        outs[i]->npoly = rep.out_npoly;
        outs[i]->limbs = rep.out_level + 1;
        optr[i] = outs[i]->ptr;
    }
    const size_t nops = rep.operands.size();
    if (rep.kind == K_SUM) {
        // one pointer table [operand][batch] per sub-batch instead of one table per operand: a sum can have 10^4 terms
        lsb::OpEnv e{ctx, *rt->ws, rt->compute, 0};
        const int n_mac = rep.aux, n_terms = (int)nops - n_mac;
        LSB_CHECK(nops + 1 <= rt->tab_cap / 2, "lsb runner: sum has too many terms for the pointer-table arena");
        const int step = (int)std::max<size_t>(1, std::min<size_t>((size_t)nb, rt->tab_cap / 2 / (nops + 1)));
        for (int b0 = 0; b0 < nb; b0 += step) {
            const int cb = std::min(step, nb - b0);
            std::vector<u64*> ptrs((nops + 1) * (size_t)cb);
            for (size_t j = 0; j < nops; j++)
                for (int i = 0; i < cb; i++) ptrs[j * cb + i] = calls[b0 + i].operands[j]->ptr;
            for (int i = 0; i < cb; i++) ptrs[nops * cb + i] = optr[b0 + i];
            u64** tab = push_table(rt, ptrs);
            lsb::op_sum_terms(e, rep.level, rep.npoly, tab, tab + (size_t)n_terms * cb, tab + nops * cb, n_mac, n_terms - n_mac, cb);
        }
        groups++;
        for (int i = 0; i < nb; i++) {
            DatumPtr d = std::make_shared<DeviceDatum>();
            d->on.resize(slots.size());
            d->on[slot] = outs[i];
            calls[i].operands.clear();
            calls[i].keepalive.clear();
            unit_done(plan.units[members[i]], d, false);
        }
        return;
    }
    std::vector<BatchPtr> tabs(nops);
    for (size_t j = 0; j < nops; j++) {
        std::vector<u64*> ptrs(nb);
        for (int i = 0; i < nb; i++) ptrs[i] = calls[i].operands[j]->ptr;
        tabs[j] = lsb::bp_table(push_table(rt, ptrs));
    }
    BatchPtr out = lsb::bp_table(push_table(rt, optr));
    lsb::OpEnv e{ctx, *rt->ws, rt->compute, 0};
    const int L = rep.level;
    switch (rep.kind) {
        case K_ADD_CC: lsb::op_add(e, L, rep.npoly, tabs[0], tabs[1], out, nb); break;
        case K_SUB_CC: lsb::op_sub(e, L, rep.npoly, tabs[0], tabs[1], out, nb); break;
        case K_ADD_CP: lsb::op_add_plain(e, L, rep.npoly, tabs[0], tabs[1], out, rep.aux != 0, nb); break;
        case K_ADD_CPR: lsb::op_add_plain_ringt(e, L, rep.npoly, tabs[0], tabs[1], out, rep.aux != 0, nb); break;
        case K_SUB_PC:  // pt - ct = -(ct - pt)
            lsb::op_add_plain(e, L, rep.npoly, tabs[0], tabs[1], out, true, nb);
            lsb::op_neg(e, L, rep.npoly, out, out, nb);
            break;
        case K_SUB_PRC:
            lsb::op_add_plain_ringt(e, L, rep.npoly, tabs[0], tabs[1], out, true, nb);
            lsb::op_neg(e, L, rep.npoly, out, out, nb);
            break;
        case K_NEG: lsb::op_neg(e, L, rep.npoly, tabs[0], out, nb); break;
        case K_MULT_CC: lsb::op_mult(e, L, tabs[0], tabs[1], out, nb); break;
        case K_MUL_CP: lsb::op_mul_plain(e, L, rep.npoly, tabs[0], tabs[1], out, nb); break;
        case K_MUL_CPR: lsb::op_mul_plain_ringt(e, L, rep.npoly, tabs[0], tabs[1], out, nb); break;
        case K_RELIN: lsb::op_relin(e, L, tabs[0], rep.key, rep.key_level, out, nb); break;
        case K_RESCALE: lsb::op_rescale(e, L, rep.npoly, tabs[0], out, nb); break;
        case K_DROP: lsb::op_drop_level(e, L, rep.npoly, tabs[0], out, nb); break;
        case K_ROTATE: lsb::op_rotate(e, L, tabs[0], rep.key, rep.key_level, rep.gal, out, nb); break;
        case K_MULT_RELIN: lsb::op_mult_relin(e, L, tabs[0], tabs[1], rep.key, rep.key_level, out, nb); break;
        case K_MULT_RELIN_RESCALE: lsb::op_mult_relin_rescale(e, L, tabs[0], tabs[1], rep.key, rep.key_level, out, nb); break;
        case K_BOOTSTRAP: {
            if (!rt->btp) rt->btp.reset(new lsb::DeviceBootstrap(ctx, btp_params(g)));
            LSB_CHECK(rt->btp->out_level() == rep.out_level, "lsb runner: bootstrap output level differs from btp_output_level");
            lsb::BootstrapStats st;
            rt->btp->run(*rt->ws, rt->compute, *rep.btp_keys, tabs[0], out, nb, &st);
            break;
        }
        case K_MAC: {
            const int cnt = rep.npoly, acc = rep.aux2;
            lsb::op_mac(e, L, cnt, tabs.data(), tabs.data() + cnt + acc, rep.aux != 0, acc != 0, acc ? tabs[cnt] : out, out, nb);
            break;
        }
        default: throw std::runtime_error("lsb runner: internal: unknown call kind");
    }
    groups++;
    for (int i = 0; i < nb; i++) {
        DatumPtr d = std::make_shared<DeviceDatum>();
        d->on.resize(slots.size());
        d->on[slot] = outs[i];
        calls[i].operands.clear();
        calls[i].keepalive.clear();
        unit_done(plan.units[members[i]], d, false);
    }
}

void RunState::flush_gpu(std::vector<int>& pending) {
    // priority first (longest remaining path), then group identical signatures
    std::stable_sort(pending.begin(), pending.end(), [&](int a, int b) { return plan.units[a].priority > plan.units[b].priority; });
    std::vector<std::string> order;
    std::unordered_map<std::string, std::pair<std::vector<int>, std::vector<GpuCall>>> groups_by_sig;
    for (int ui : pending) {
        const Unit& u = plan.units[ui];
        try {
            if (failed.load()) throw std::runtime_error("cancelled");
            const int slot = slot_of(u);
            auto in = gather_inputs(u);
            GpuCall c = describe(u, in, slot);
            std::string sig = c.signature(slot);
            auto it = groups_by_sig.find(sig);
            if (it == groups_by_sig.end()) {
                order.push_back(sig);
                it = groups_by_sig.emplace(sig, std::make_pair(std::vector<int>(), std::vector<GpuCall>())).first;
            }
            it->second.first.push_back(ui);
            it->second.second.push_back(std::move(c));
        } catch (const std::exception& e) {
            if (std::string(e.what()) != "cancelled") fail("node '" + u.nodes[0]->id + "': " + e.what());
            unit_done(u, std::any(), false);
        }
    }
    pending.clear();
    auto run_sigs = [&](const std::vector<const std::string*>& sigs) {
        for (const std::string* sp : sigs) {
            auto& grp = groups_by_sig.at(*sp);
            const int slot = slot_of(plan.units[grp.first[0]]);
            try {
                if (failed.load()) throw std::runtime_error("cancelled");
                GpuCall rep = grp.second[0];
                exec_group(slot, grp.first, rep, grp.second);
            } catch (const std::exception& e) {
                if (std::string(e.what()) != "cancelled") fail(std::string("GPU batch (") + plan.units[grp.first[0]].nodes[0]->id + ", " +
                                                               std::to_string(grp.first.size()) + " nodes): " + e.what());
                // publish empty outputs so that the graph drains
                for (int ui : grp.first) {
                    bool already;
                    {
                        std::lock_guard<std::mutex> lk(mu);
                        already = avail.count(plan.units[ui].output->index) != 0;
                    }
                    if (!already) unit_done(plan.units[ui], std::any(), false);
                }
            }
        }
    };
    std::vector<const std::string*> all;
    for (const std::string& sig : order) all.push_back(&sig);
    run_sigs(all);
}


// ---------------------------------------------------------------------------------------------
// dispatcher (runs on the caller's thread; replaces run_tasks, cpu_task_utils.h:377-569, and the
// submit_gpu_task lambda of gpu_wrapper.cu:204-352)
// ---------------------------------------------------------------------------------------------
void RunState::dispatch(int q) {
    std::vector<int>& ready = ready_q[q];
    const size_t total_units = plan.units.size();
    const size_t batch_min = env_mb("LSB_BATCH_MIN", 128);
    const auto linger = std::chrono::microseconds(env_mb("LSB_BATCH_LINGER_US", 1000));
    std::vector<int> gpu_pending;
    const bool trace = getenv("LSB_TRACE") != nullptr;  // dispatcher timeline on stderr
    const auto t_start = std::chrono::steady_clock::now();
    auto pending_since = std::chrono::steady_clock::now();
    for (;;) {
        std::vector<int> batch;
        int flying;
        {
            std::unique_lock<std::mutex> lk(mu);
            if (ready.empty() && gpu_pending.empty()) {
                if (done_units >= total_units || aborted) break;
                cv.wait(lk, [&] { return !ready.empty() || done_units >= total_units || aborted; });
                if (ready.empty()) break;
            } else if (ready.empty()) {
                // GPU work is waiting to be batched: give in-flight host nodes a moment to add to it
                if (feeding > 0 && gpu_pending.size() < batch_min) cv.wait_until(lk, pending_since + linger, [&] { return !ready.empty(); });
            }
            batch.swap(ready);
            flying = feeding;
        }
        std::stable_sort(batch.begin(), batch.end(), [&](int a, int b) { return plan.units[a].priority > plan.units[b].priority; });
        for (int ui : batch) {
            const Unit& u = plan.units[ui];
            switch (u.kind) {
                case UnitKind::HOST:
                case UnitKind::LOAD: {
                    {
                        std::lock_guard<std::mutex> lk(mu);
                        in_flight++;
                        if (feeds_gpu(u)) feeding++;
                    }
                    const Unit* up = &u;
                    if (u.kind == UnitKind::HOST) {
                        const bool drain = op_of(u.nodes[0]) == OperationType::IMPORT_FROM_ABI;
                        pool.submit((drain ? kDrainPriority : 0) + u.priority, [this, up] { run_host_unit(*up); });
                    } else {
                        pool.submit(kLoadPriority - std::min(u.order_key, kLoadPriority - 1), [this, up] { run_load_unit(*up); });
                    }
                    break;
                }
                case UnitKind::STORE: start_store_unit(u); break;
                case UnitKind::GPU:
                    if (gpu_pending.empty()) pending_since = std::chrono::steady_clock::now();
                    gpu_pending.push_back(ui);
                    break;
            }
        }
        if (!gpu_pending.empty()) {
            bool more_ready;
            {
                std::lock_guard<std::mutex> lk(mu);
                more_ready = !ready.empty();
                flying = feeding;
            }
            const bool waited = std::chrono::steady_clock::now() - pending_since >= linger;
            if (gpu_pending.size() >= batch_min || (!more_ready && (flying == 0 || waited))) {
                const auto f0 = std::chrono::steady_clock::now();
                const size_t np = gpu_pending.size();
                flush_gpu(gpu_pending);
                if (trace) {
                    const auto f1 = std::chrono::steady_clock::now();
                    fprintf(stderr, "[lsb trace] q%d +%8.3f ms flush %4zu units in %7.3f ms (feeding %d)\n",
                            q, std::chrono::duration<double, std::milli>(f0 - t_start).count(), np,
                            std::chrono::duration<double, std::milli>(f1 - f0).count(), flying);
                }
            }
        }
    }
    if (trace)
        fprintf(stderr, "[lsb trace] q%d dispatch done at +%.3f ms\n", q,
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count());
}

static int run_task(GpuTask& task, CArgument* input_args, uint64_t n_in, CArgument* output_args, uint64_t n_out,
                    progress_callback_t cb, void* ud, int gpu_device) {
    std::lock_guard<std::mutex> run_lock(task.run_mu);
    TaskGraph& g = *task.graph;
    std::vector<int> devs = task.devices.empty() ? std::vector<int>{gpu_device} : task.devices;
    devs.resize(worthwhile_slots(task.plan, (int)devs.size(), (double)g.param.n));
    const int hw = (int)std::thread::hardware_concurrency();
    // worker threads: the copies they make are bound by host DRAM bandwidth, not by cores -- measured on a 16-core B200
    // host (cfg2 x1024 through run_fhe_gpu_task, profiles/r02_e2e_host_path.txt): 8 threads 62 ms, 12 threads 65 ms,
    // 15 threads 73 ms per call -- so the pool is capped at 8 (the reference's pool cap is min(32, cores),
    // cpu_wrapper.cpp:57) and shared fairly when several ranks run on one node (torchrun exports LOCAL_WORLD_SIZE);
    // LSB_HOST_THREADS overrides
    int share = hw > 0 ? hw : 4;
    if (const char* lw = getenv("LOCAL_WORLD_SIZE")) share = std::max(2, share / std::max(1, atoi(lw)));
    int nthreads = std::max(2, std::min(9, share) - 1);
    if (const char* ht = getenv("LSB_HOST_THREADS")) nthreads = std::max(1, atoi(ht));
    static WorkerPool pool(nthreads);  // shared by all tasks of the process
    RunState rs(task, pool, cb, ud);
    task.stat_h2d = task.stat_d2h = task.stat_zero_copy = task.stat_key_cached = 0;
    task.ns_export = task.ns_stage = task.ns_ring_wait = task.ns_d2h_wait = task.ns_import = 0;
    const long long t_run0 = now_ns();
    long long launches0 = 0;
    {
        std::map<int, int> seen;
        for (int d : devs) {
            rs.slots.push_back(task.runtime(d, seen[d]++));
            launches0 += rs.slots.back()->ctx->prof.launches.load();
        }
    }
    rs.slot_of_component = assign_components(task.plan, (int)devs.size());
    if (task.ctx_factory) rs.frontend_ctx = task.ctx_factory(g.algo, json_dump(g.parameter));

    // inputs: flattened handles in mega_ag.inputs order; all glk data share one handle (cpu_task_utils.h:295-319)
    std::vector<void*> handles;
    for (uint64_t i = 0; i < n_in; i++) {
        void** arr = static_cast<void**>(input_args[i].data);
        for (int j = 0; j < input_args[i].size; j++) handles.push_back(arr[j]);
    }
    {
        std::shared_ptr<void> shared_glk;
        size_t hi = 0;
        for (NodeIndex idx : g.inputs) {
            const DatumNode& d = g.data.at(idx);
            if (d.fhe_prop.has_value() && d.datum_type == TYPE_GALOIS_KEY) {
                if (!shared_glk) {
                    LSB_CHECK(hi < handles.size(), "lsb runner: fewer input handles than graph inputs");
                    shared_glk = std::shared_ptr<void>(handles[hi++], [](void*) {});
                }
                rs.avail[idx] = shared_glk;
            } else {
                LSB_CHECK(hi < handles.size(), "lsb runner: fewer input handles than graph inputs");
                rs.avail[idx] = std::shared_ptr<void>(handles[hi++], [](void*) {});
            }
        }
    }
    {
        size_t oi = 0;
        for (uint64_t i = 0; i < n_out; i++) {
            void** arr = static_cast<void**>(output_args[i].data);
            for (int j = 0; j < output_args[i].size; j++) {
                LSB_CHECK(oi < g.outputs.size(), "lsb runner: more output handles than graph outputs");
                rs.output_handles[g.outputs[oi++]] = arr[j];
            }
        }
    }
    rs.ready_q.assign(devs.size(), {});
    rs.remaining.assign(task.plan.units.size(), 0);
    for (const Unit& u : task.plan.units) {
        int missing = 0;
        for (DatumNode* in : u.ext_inputs)
            if (!rs.avail.count(in->index)) missing++;
        rs.remaining[u.id] = missing;
        if (missing == 0) rs.ready_q[rs.queue_of(u.id)].push_back(u.id);
    }
    for (auto& kv : task.plan.consumers) rs.refcnt[kv.first] = (int)kv.second.size();

    // one dispatcher per device (each GPU has its own scheduler): components never span devices, so the
    // queues are independent; the caller's thread serves device slot 0 and the host-only units
    {
        auto serve = [&rs](int q) {
            try {
                rs.dispatch(q);
            } catch (const std::exception& e) {
                rs.fail(std::string("dispatcher: ") + e.what());
                {
                    std::lock_guard<std::mutex> lk(rs.mu);
                    rs.aborted = true;
                }
                rs.cv.notify_all();
            }
        };
        std::vector<std::thread> others;
        for (int q = 1; q < (int)devs.size(); q++) others.emplace_back(serve, q);
        serve(0);
        for (auto& t : others) t.join();
    }
    {
        // tasks already handed to the process-wide worker pool capture `rs`: on an aborted run the dispatchers return
        // early, so wait here until every one of them has finished (they return quickly once `failed` is set)
        std::unique_lock<std::mutex> lk(rs.mu);
        rs.cv.wait(lk, [&] { return rs.in_flight == 0; });
    }

    for (DeviceRuntime* rt : rs.slots) {
        LSB_CUDA(cudaSetDevice(rt->device));
        LSB_CUDA(cudaStreamSynchronize(rt->compute));
        LSB_CUDA(cudaStreamSynchronize(rt->h2d));
        for (auto st : rt->h2d_extra) LSB_CUDA(cudaStreamSynchronize(st));
        LSB_CUDA(cudaStreamSynchronize(rt->d2h));
    }
    rs.avail.clear();
    long long launches1 = 0;
    for (DeviceRuntime* rt : rs.slots) launches1 += rt->ctx->prof.launches.load();
    task.ns_wall = now_ns() - t_run0;
    task.stat_launches = launches1 - launches0;
    task.stat_groups = rs.groups;
    task.stat_fused = task.plan.fused_nodes;
    if (rs.failed.load()) {
        g_runner_error = rs.error;
        return 1;
    }
    return 0;
}

}  // namespace lsr

// ---------------------------------------------------------------------------------------------
// C ABI (mega_ag_runners/wrapper.h:67-85, gpu_wrapper.cu:481-530)
// ---------------------------------------------------------------------------------------------
using lsr::GpuTask;

void bind_gpu_executor(ComputeNode& node, Algo algorithm) {
    if (!node.fhe_prop.has_value()) throw std::runtime_error("FHE property not found for compute node");
    const OperationType op = node.fhe_prop->op_type;
    bool ok = false;
    switch (op) {
        case OperationType::ADD:
        case OperationType::SUB:
        case OperationType::NEGATE:
        case OperationType::MULTIPLY:
        case OperationType::RELINEARIZE:
        case OperationType::RESCALE:
        case OperationType::ROTATE_COL:
        case OperationType::ROTATE_ROW:
        case OperationType::MAC_W_PARTIAL_SUM:
        case OperationType::MAC_WO_PARTIAL_SUM: ok = algorithm == ALGO_BFV || algorithm == ALGO_CKKS; break;
        case OperationType::DROP_LEVEL:
        case OperationType::BOOTSTRAP: ok = algorithm == ALGO_CKKS; break;
        default: ok = false;
    }
    if (!ok) throw std::runtime_error(algorithm == ALGO_BFV ? "Unsupported operation type for GPU BFV" : "Unsupported operation type for GPU CKKS");
    const std::string id = node.id;
    node.executor = [id](ExecutionContext&, const std::unordered_map<NodeIndex, std::any>&, std::any&, const ComputeNode&) {
        throw std::runtime_error("lsb runner: node '" + id + "' is executed by the batched dispatcher of run_fhe_gpu_task "
                                 "(create_fhe_gpu_task / run_fhe_gpu_task), not by a per-node executor");
    };
}

extern "C" {

fhe_task_handle create_fhe_gpu_task(const char* project_path) {
    if (!project_path) throw std::runtime_error("create_fhe_gpu_task: null project path");
    std::unique_ptr<GpuTask> t(new GpuTask());
    t->graph.reset(lsr::TaskGraph::load(std::string(project_path) + "/mega_ag.json"));
    if (const char* kc = getenv("LSB_KEY_CACHE")) t->key_cache_on = atoi(kc) != 0;
    const char* nofuse = getenv("LSB_NO_FUSE");
    const char* nohost = getenv("LSB_NO_HOST_FUSE");
    t->plan = lsr::build_plan(*t->graph, !(nofuse && atoi(nofuse) != 0), !(nohost && atoi(nohost) != 0));
    return reinterpret_cast<fhe_task_handle>(t.release());
}

void release_fhe_gpu_task(fhe_task_handle handle) { delete reinterpret_cast<GpuTask*>(handle); }

void bind_gpu_task_abi_bridge_executors(fhe_task_handle handle, void* abi_export_executor, void* abi_import_executor) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t || !abi_export_executor || !abi_import_executor) throw std::runtime_error("bind_gpu_task_abi_bridge_executors: null argument");
    t->graph->bind_abi_bridge_executors(*reinterpret_cast<ExecutorFunc*>(abi_export_executor),
                                        *reinterpret_cast<ExecutorFunc*>(abi_import_executor));
}

void bind_gpu_task_custom_executors(fhe_task_handle handle, const char** custom_types, void** executors, uint64_t n_executors) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t) throw std::runtime_error("bind_gpu_task_custom_executors: null task");
    std::unordered_map<std::string, ExecutorFunc> m;
    for (uint64_t i = 0; i < n_executors; i++) m[std::string(custom_types[i])] = *reinterpret_cast<ExecutorFunc*>(executors[i]);
    t->graph->bind_custom_executors(m);
}

int run_fhe_gpu_task(fhe_task_handle handle, CArgument* input_args, uint64_t n_in_args, CArgument* output_args,
                     uint64_t n_out_args, progress_callback_t progress_cb, void* user_data, int gpu_device) {
    lsr::g_runner_error.clear();
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    try {
        if (!t) throw std::runtime_error("run_fhe_gpu_task: null task");
        return lsr::run_task(*t, input_args, n_in_args, output_args, n_out_args, progress_cb, user_data, gpu_device);
    } catch (const std::exception& e) {
        lsr::g_runner_error = e.what();
        return 1;
    } catch (...) {
        lsr::g_runner_error = "run_fhe_gpu_task: unknown exception";
        return 1;
    }
}

void lsb_gpu_task_set_devices(fhe_task_handle handle, const int* devices, int n_devices) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t) return;
    std::lock_guard<std::mutex> lk(t->run_mu);
    t->devices.assign(devices, devices + (n_devices > 0 ? n_devices : 0));
    std::lock_guard<std::mutex> lk2(t->key_mu);
    t->key_cache.clear();  // cached keys live on the device slots of the previous configuration
}

void lsb_gpu_task_set_key_cache(fhe_task_handle handle, int on) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t) return;
    std::lock_guard<std::mutex> lk(t->run_mu);
    t->key_cache_on = on != 0;
    if (!on) {
        std::lock_guard<std::mutex> lk2(t->key_mu);
        t->key_cache.clear();
    }
}
void lsb_gpu_task_invalidate_keys(fhe_task_handle handle) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t) return;
    std::lock_guard<std::mutex> lk(t->run_mu);
    std::lock_guard<std::mutex> lk2(t->key_mu);
    t->key_cache.clear();
}
void lsb_gpu_task_transfer_stats(fhe_task_handle handle, long long* h2d_bytes, long long* d2h_bytes, long long* zero_copy_bytes,
                                 long long* key_bytes_cached) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t) return;
    if (h2d_bytes) *h2d_bytes = t->stat_h2d.load();
    if (d2h_bytes) *d2h_bytes = t->stat_d2h.load();
    if (zero_copy_bytes) *zero_copy_bytes = t->stat_zero_copy.load();
    if (key_bytes_cached) *key_bytes_cached = t->stat_key_cached.load();
}

void lsb_gpu_task_host_profile(fhe_task_handle handle, double* ms6) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t || !ms6) return;
    ms6[0] = 1e-6 * (double)t->ns_wall.load();
    ms6[1] = 1e-6 * (double)t->ns_export.load();
    ms6[2] = 1e-6 * (double)t->ns_stage.load();
    ms6[3] = 1e-6 * (double)t->ns_ring_wait.load();
    ms6[4] = 1e-6 * (double)t->ns_d2h_wait.load();
    ms6[5] = 1e-6 * (double)t->ns_import.load();
}

void lsb_gpu_task_last_stats(fhe_task_handle handle, long long* kernel_launches, long long* batched_groups, long long* fused_nodes) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t) return;
    if (kernel_launches) *kernel_launches = t->stat_launches;
    if (batched_groups) *batched_groups = t->stat_groups;
    if (fused_nodes) *fused_nodes = t->stat_fused;
}

void lsb_gpu_task_set_frontend_context_factory(fhe_task_handle handle, void* factory) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t) return;
    t->ctx_factory = factory ? *reinterpret_cast<FrontendContextFactory*>(factory) : FrontendContextFactory();
}

const char* lsb_gpu_task_last_error(void) { return lsr::g_runner_error.c_str(); }

// graph introspection for host-side tests (no GPU needed): counts after bridge insertion and planning
int lsb_gpu_task_describe(fhe_task_handle handle, long long* out8) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t || !out8) return 1;
    long long host = 0, load = 0, store = 0, gpu = 0;
    for (auto& u : t->plan.units) {
        if (u.kind == lsr::UnitKind::HOST) host++;
        else if (u.kind == lsr::UnitKind::LOAD) load++;
        else if (u.kind == lsr::UnitKind::STORE) store++;
        else gpu++;
    }
    out8[0] = (long long)t->graph->computes.size();
    out8[1] = (long long)t->graph->data.size();
    out8[2] = host;
    out8[3] = load;
    out8[4] = store;
    out8[5] = gpu;
    out8[6] = t->plan.fused_nodes;
    out8[7] = t->plan.n_components;
    return 0;
}

// component -> device slot assignment for `n_slots` devices (host-side test of the sharding)
int lsb_gpu_task_shard_preview(fhe_task_handle handle, int n_slots, int* slot_of_component, int capacity) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t) return -1;
    std::vector<int> s = lsr::assign_components(t->plan, n_slots);
    for (int i = 0; i < (int)s.size() && i < capacity; i++) slot_of_component[i] = s[i];
    return (int)s.size();
}

// how many of `n_devices` the next run would open (lsb_gpu_task_set_devices is a request: small graphs use fewer)
int lsb_gpu_task_worthwhile_devices(fhe_task_handle handle, int n_devices) {
    GpuTask* t = reinterpret_cast<GpuTask*>(handle);
    if (!t) return -1;
    return lsr::worthwhile_slots(t->plan, n_devices, (double)t->graph->param.n);
}

}  // extern "C"
