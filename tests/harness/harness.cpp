// Test plug-in for libgpu_mega_ag_runner.so (TEST INFRASTRUCTURE).
//
// Plays the role the reference's SDK (cxx_sdk_v2/cxx_fhe_task_gpu.cpp:32-117) or a plug-in
// (plug-in/SEAL, plug-in/lattigo) plays: it owns the "frontend handles", binds EXPORT_TO_ABI /
// IMPORT_FROM_ABI executors through bind_gpu_task_abi_bridge_executors and drives
// create_/run_/release_fhe_gpu_task with CArgument arrays — exactly the calls FheTaskGpu makes,
// without needing liblattigo.  Handles are plain limb buffers owned by the Python test.
#include <any>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "lattisense_gpu_runner.h"
#include "lattisense_mega_ag.hpp"

extern "C" const char* lsb_gpu_task_last_error(void);

extern "C" {
// one frontend handle (mirrors what BfvCiphertext / CkksPlaintext / RelinKey / GaloisKey handles carry)
struct THandle {
    int kind;     // 0 ciphertext, 1 plaintext (pt / pt_mul: level+1 limbs), 2 pt_ringt (1 limb), 3 key-switch key, 4 Galois key set,
                  // 5 custom payload (a message: `level+1` rows of n words, what a CustomData handle carries)
    int level;    // ciphertext/plaintext level, or key level
    int degree;   // ciphertext degree
    int n;        // ring degree
    int k;        // keys: number of special primes
    uint64_t* data;  // ct: [degree+1][level+1][n]; pt: [limbs][n]; key: [beta][2][level+1+k][n]
    int n_keys;                // kind 4
    uint64_t* galois_elements; // kind 4
    THandle** keys;            // kind 4: key-switch key handles
};
}

namespace {

thread_local std::string g_err;

THandle* handle_of(const DatumNode* d, const std::any& a, std::shared_ptr<THandle>* owner) {
    if (d->is_input) return static_cast<THandle*>(std::any_cast<std::shared_ptr<void>>(a).get());
    *owner = std::any_cast<std::shared_ptr<THandle>>(a);
    return owner->get();
}

void fill_poly(CPolynomial& p, const uint64_t* src, int n) {
    for (int i = 0; i < p.n_component; i++) lsb_host_copy(p.components[i].data, src + (size_t)i * n, (size_t)n * 8);
}

// key handle at key level KL -> CKeySwitchKey at `level` (plug-in/lattigo/acc/c_struct_import_export.go:83-135):
// beta' = ceil((level+1)/k) digits, Q limbs 0..level then the k P limbs
void export_ksk(const THandle* h, int level, CKeySwitchKey* out) {
    const int KL = h->level, k = h->k, n = h->n;
    if (level < 0 || level > KL) throw std::runtime_error("harness: key requested above its level");
    const int beta = (level + 1 + k - 1) / k;
    const int src_limbs = KL + 1 + k, dst_limbs = level + 1 + k;
    out->n_public_key = beta;
    out->public_keys = (CPublicKey*)malloc(sizeof(CPublicKey) * beta);
    for (int d = 0; d < beta; d++) {
        CPublicKey& pk = out->public_keys[d];
        pk.level = level;
        pk.degree = 1;
        pk.polys = (CPolynomial*)malloc(sizeof(CPolynomial) * 2);
        for (int j = 0; j < 2; j++) {
            CPolynomial& poly = pk.polys[j];
            poly.n_component = dst_limbs;
            poly.components = (CComponent*)malloc(sizeof(CComponent) * dst_limbs);
            const uint64_t* base = h->data + ((size_t)(d * 2 + j) * src_limbs) * n;
            for (int t = 0; t < dst_limbs; t++) {
                const int s = t <= level ? t : KL + 1 + (t - level - 1);
                alloc_component(&poly.components[t], n);
                lsb_host_copy(poly.components[t].data, base + (size_t)s * n, (size_t)n * 8);
            }
        }
    }
}

ExecutorFunc make_export() {
    return [](ExecutionContext&, const std::unordered_map<NodeIndex, std::any>& inputs, std::any& output, const ComputeNode& self) {
        const DatumNode* d = self.input_nodes[0];
        std::shared_ptr<THandle> owner;
        THandle* h = handle_of(d, inputs.at(d->index), &owner);
        if (!h) throw std::runtime_error("harness: null handle for " + d->id);
        const int level = d->fhe_prop.has_value() ? d->fhe_prop->level : -1;
        switch (d->datum_type) {
            case TYPE_CIPHERTEXT: {
                if (h->kind != 0 || h->level != level) throw std::runtime_error("harness: ciphertext handle does not match node " + d->id);
                CCiphertext* c = (CCiphertext*)malloc(sizeof(CCiphertext));
                alloc_ciphertext(c, h->degree, h->level, h->n);
                for (int p = 0; p <= h->degree; p++) fill_poly(c->polys[p], h->data + (size_t)p * (h->level + 1) * h->n, h->n);
                output = std::shared_ptr<CCiphertext>(c, [](CCiphertext* x) {
                    free_ciphertext(x);
                    free(x);
                });
                break;
            }
            case TYPE_PLAINTEXT: {
                const bool ringt = d->fhe_prop->p.has_value() && d->fhe_prop->p->is_ringt;
                if (h->kind != (ringt ? 2 : 1)) throw std::runtime_error("harness: plaintext handle kind does not match node " + d->id);
                const int lv = ringt ? 0 : h->level;
                CPlaintext* c = (CPlaintext*)malloc(sizeof(CPlaintext));
                alloc_plaintext(c, lv, h->n);
                fill_poly(c->poly, h->data, h->n);
                output = std::shared_ptr<CPlaintext>(c, [](CPlaintext* x) {
                    free_plaintext(x);
                    free(x);
                });
                break;
            }
            case TYPE_RELIN_KEY:
            case TYPE_SWITCH_KEY: {
                if (h->kind != 3) throw std::runtime_error("harness: key handle expected for node " + d->id);
                CKeySwitchKey* c = (CKeySwitchKey*)malloc(sizeof(CKeySwitchKey));
                export_ksk(h, level, c);
                auto del = [](CKeySwitchKey* x) {
                    free_relin_key(x);
                    free(x);
                };
                if (d->datum_type == TYPE_RELIN_KEY) output = std::shared_ptr<CRelinKey>(c, del);
                else output = std::shared_ptr<CKeySwitchKey>(c, del);
                break;
            }
            case TYPE_GALOIS_KEY: {
                if (h->kind != 4) throw std::runtime_error("harness: Galois key set expected for node " + d->id);
                uint64_t g = d->fhe_prop->p.has_value() ? d->fhe_prop->p->galois_element : 0;
                const THandle* key = nullptr;
                for (int i = 0; i < h->n_keys; i++)
                    if (h->galois_elements[i] == g) key = h->keys[i];
                if (!key) throw std::runtime_error("harness: no Galois key for element " + std::to_string(g));
                CGaloisKey* c = (CGaloisKey*)malloc(sizeof(CGaloisKey));
                set_galois_key_steps(c, &g, 1);
                c->key_switch_keys = (CKeySwitchKey*)malloc(sizeof(CKeySwitchKey));
                export_ksk(key, level, &c->key_switch_keys[0]);
                output = std::shared_ptr<CGaloisKey>(c, [](CGaloisKey* x) {
                    free_galois_key(x);
                    free(x);
                });
                break;
            }
            case TYPE_CUSTOM:  // custom data input: handed on as it is (mega_ag.cpp:424-434, "Case 0"; the SDK returns shared_ptr<CustomData>)
                if (h->kind != 5) throw std::runtime_error("harness: custom payload handle expected for node " + d->id);
                output = std::shared_ptr<THandle>(h, [](THandle*) {});
                break;
            default: throw std::runtime_error("harness: unsupported data type for EXPORT_TO_ABI");
        }
    };
}

void copy_ct(THandle* dst, const CCiphertext* c) {
    if (dst->kind != 0 || dst->level != c->level || dst->degree != c->degree)
        throw std::runtime_error("harness: output handle shape does not match the produced ciphertext");
    for (int p = 0; p <= c->degree; p++)
        for (int i = 0; i <= c->level; i++)
            lsb_host_copy(dst->data + ((size_t)p * (c->level + 1) + i) * dst->n, c->polys[p].components[i].data, (size_t)dst->n * 8);
}

ExecutorFunc make_import() {
    return [](ExecutionContext& ctx, const std::unordered_map<NodeIndex, std::any>& inputs, std::any& output, const ComputeNode& self) {
        const DatumNode* d = self.input_nodes[0];
        if (d->datum_type != TYPE_CIPHERTEXT) throw std::runtime_error("harness: unsupported data type for IMPORT_FROM_ABI");
        const std::any& in = inputs.at(d->index);
        if (in.type() == typeid(std::shared_ptr<CCiphertext>)) {
            auto c = std::any_cast<std::shared_ptr<CCiphertext>>(in);
            if (!ctx.other_args.empty()) {
                THandle* dst = static_cast<THandle*>(ctx.get_other_arg<void>(0));
                copy_ct(dst, c.get());
                output = std::shared_ptr<void>(dst, [](void*) {});
            } else {  // intermediate datum consumed by a custom node: new handle owning its limbs
                const int n = c->polys[0].components[0].n;
                THandle* h = new THandle();
                std::memset(h, 0, sizeof(THandle));
                h->kind = 0;
                h->level = c->level;
                h->degree = c->degree;
                h->n = n;
                h->data = (uint64_t*)malloc((size_t)(c->degree + 1) * (c->level + 1) * n * 8);
                copy_ct(h, c.get());
                output = std::shared_ptr<THandle>(h, [](THandle* x) {
                    free(x->data);
                    delete x;
                });
            }
        } else {  // handle produced by a custom node that is a graph output
            auto src = std::any_cast<std::shared_ptr<THandle>>(in);
            THandle* dst = static_cast<THandle*>(ctx.get_other_arg<void>(0));
            if (!dst) throw std::runtime_error("harness: handle IMPORT_FROM_ABI requires a pre-allocated destination");
            std::memcpy(dst->data, src->data, (size_t)(src->degree + 1) * (src->level + 1) * src->n * 8);
            output = std::shared_ptr<void>(dst, [](void*) {});
        }
    };
}

// custom node "passthrough": hands its input handle on; "add_one_limbwise": out = in + 1 on every residue
// (shows that the value really went through the host)
std::atomic<int> g_custom_calls{0};
ExecutorFunc make_passthrough(bool add_one) {
    return [add_one](ExecutionContext&, const std::unordered_map<NodeIndex, std::any>& inputs, std::any& output, const ComputeNode& self) {
        const DatumNode* d = self.input_nodes[0];
        auto src = std::any_cast<std::shared_ptr<THandle>>(inputs.at(d->index));
        g_custom_calls++;
        if (!add_one) {
            output = src;
            return;
        }
        throw std::runtime_error("harness: add_one custom node used without moduli");
    };
}

// custom nodes "encode_ringt" / "encode" (unittests/test_gpu_bfv.cpp:1109-1131): a custom payload (kind 5) becomes a
// pt_ringt / plaintext handle the device consumes through EXPORT_TO_ABI; the payload is taken as already encoded (the
// test plug-in has no encoder), `encode` reads its level from the node's attributes like the reference executor does
ExecutorFunc make_encode(bool ringt) {
    return [ringt](ExecutionContext&, const std::unordered_map<NodeIndex, std::any>& inputs, std::any& output, const ComputeNode& self) {
        const DatumNode* d = self.input_nodes[0];
        std::shared_ptr<THandle> owner;
        THandle* src = handle_of(d, inputs.at(d->index), &owner);
        if (!src || src->kind != 5) throw std::runtime_error("harness: encode expects a custom payload handle");
        int level = 0;
        if (!ringt) {
            if (!self.custom_prop.has_value()) throw std::runtime_error("harness: custom property not found for encode");
#ifdef LATTISENSE_WITH_NLOHMANN_JSON
            level = self.custom_prop->attributes.at("level").get<int>();
#else
            level = self.custom_prop->attributes.at("level").as_int();
#endif
        }
        if (src->level != level) throw std::runtime_error("harness: payload rows do not match the encode level");
        g_custom_calls++;
        THandle* h = new THandle(*src);
        h->kind = ringt ? 2 : 1;
        h->level = level;
        output = std::shared_ptr<THandle>(h);  // non-owning view of the payload's words
    };
}

struct Progress {
    std::atomic<int> calls{0};
    std::atomic<int> last_completed{0};
    std::atomic<int> total{0};
};
void on_progress(int completed, int total, void* ud) {
    Progress* p = static_cast<Progress*>(ud);
    p->calls++;
    p->last_completed = completed;
    p->total = total;
}

}  // namespace

extern "C" {

const char* th_last_error() { return g_err.c_str(); }

void* th_open(const char* project_path, int bind_executors) {
    g_err.clear();
    try {
        fhe_task_handle t = create_fhe_gpu_task(project_path);
        if (bind_executors) {
            // executors are copied by the callee: locals may die right after the call (plug-in/lattigo/acc/gpu_runner.go:94-96)
            ExecutorFunc ex = make_export(), im = make_import();
            bind_gpu_task_abi_bridge_executors(t, &ex, &im);
            ExecutorFunc pass = make_passthrough(false), er = make_encode(true), en = make_encode(false);
            const char* types[3] = {"passthrough", "encode_ringt", "encode"};
            void* fns[3] = {&pass, &er, &en};
            bind_gpu_task_custom_executors(t, types, fns, 3);
        }
        return t;
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}

void th_close(void* t) { release_fhe_gpu_task((fhe_task_handle)t); }

// args: per argument its type, level and size; handles are flattened in argument order
int th_run(void* task, int n_in, const int* in_types, const int* in_levels, const int* in_sizes, THandle** in_handles, int n_out,
           const int* out_types, const int* out_levels, const int* out_sizes, THandle** out_handles, int device, int* progress_out3) {
    g_err.clear();
    std::vector<CArgument> in(n_in), out(n_out);
    std::vector<std::string> names;
    names.reserve(n_in + n_out);
    size_t off = 0;
    for (int i = 0; i < n_in; i++) {
        names.push_back("in_" + std::to_string(i));
        in[i] = CArgument{names.back().c_str(), (DataType)in_types[i], (void*)(in_handles + off), in_levels[i], in_sizes[i]};
        off += in_sizes[i];
    }
    off = 0;
    for (int i = 0; i < n_out; i++) {
        names.push_back("out_" + std::to_string(i));
        out[i] = CArgument{names.back().c_str(), (DataType)out_types[i], (void*)(out_handles + off), out_levels[i], out_sizes[i]};
        off += out_sizes[i];
    }
    Progress prog;
    int rc;
    try {
        rc = run_fhe_gpu_task((fhe_task_handle)task, in.data(), n_in, out.data(), n_out, on_progress, &prog, device);
    } catch (const std::exception& e) {
        g_err = e.what();
        return 2;
    }
    if (rc != 0) g_err = lsb_gpu_task_last_error();
    if (progress_out3) {
        progress_out3[0] = prog.calls;
        progress_out3[1] = prog.last_completed;
        progress_out3[2] = prog.total;
    }
    return rc;
}

int th_custom_calls() { return g_custom_calls.load(); }

}  // extern "C"
