// Allocation helpers for the ABI structs (same ownership rules as the reference's abi/c_structs.c:23-98:
// one malloc per limb, freed limb by limb).  The Lattigo import/export shims of that file are not part of
// this library: they belong to the caller that links liblattigo.
//
// Zero-copy EXPORT: while the runner executes a bound EXPORT_TO_ABI executor on one of its
// worker threads it lends that thread a region of the pinned upload ring (lsr_export_arena_begin); alloc_component then
// hands out consecutive limbs of that region instead of malloc'd blocks, so the executor writes the residues straight
// into page-locked memory and the runner's H2D copy starts from there -- no second host copy, no page faults on fresh
// 128 KiB mallocs.  free_* recognise pinned pointers by address range and leave them to the ring.
#include <stdlib.h>
#include <atomic>
#include <mutex>

#include "lattisense_gpu_runner.h"
#include "lsr_abi_arena.h"

namespace {
struct Range {
    std::atomic<uintptr_t> begin{0}, end{0};
};
constexpr int kMaxRanges = 64;
Range g_ranges[kMaxRanges];
std::mutex g_range_mu;
thread_local lsr::ExportArena* t_arena = nullptr;

inline bool in_pinned(const void* p) {
    const uintptr_t a = (uintptr_t)p;
    for (int i = 0; i < kMaxRanges; i++) {
        const uintptr_t b = g_ranges[i].begin.load(std::memory_order_acquire);
        if (b && a >= b && a < g_ranges[i].end.load(std::memory_order_acquire)) return true;
    }
    return false;
}
inline void free_limb(uint64_t* p) {
    if (p && !in_pinned(p)) free(p);
}
}  // namespace

namespace lsr {
void register_pinned_range(const void* base, size_t bytes) {
    std::lock_guard<std::mutex> lk(g_range_mu);
    for (int i = 0; i < kMaxRanges; i++)
        if (g_ranges[i].begin.load() == 0) {
            g_ranges[i].end.store((uintptr_t)base + bytes, std::memory_order_release);
            g_ranges[i].begin.store((uintptr_t)base, std::memory_order_release);
            return;
        }
}
void unregister_pinned_range(const void* base) {
    std::lock_guard<std::mutex> lk(g_range_mu);
    for (int i = 0; i < kMaxRanges; i++)
        if (g_ranges[i].begin.load() == (uintptr_t)base) g_ranges[i].begin.store(0, std::memory_order_release);
}
void export_arena_begin(ExportArena* a) { t_arena = a; }
void export_arena_end() { t_arena = nullptr; }
}  // namespace lsr

extern "C" {

void alloc_component(CComponent* component, int n) {
    component->n = n;
    const size_t bytes = (size_t)n * sizeof(uint64_t);
    lsr::ExportArena* a = t_arena;
    if (a && a->used + bytes <= a->cap) {
        component->data = (uint64_t*)(a->base + a->used);
        a->used += bytes;
        return;
    }
    if (a) a->overflow = true;
    component->data = (uint64_t*)malloc(bytes);
}

static void alloc_poly(CPolynomial* poly, int n_component, int n) {
    poly->n_component = n_component;
    poly->components = (CComponent*)malloc((size_t)n_component * sizeof(CComponent));
    for (int i = 0; i < n_component; i++) alloc_component(&poly->components[i], n);
}
static void free_poly(CPolynomial* poly) {
    for (int i = 0; i < poly->n_component; i++) free_limb(poly->components[i].data);
    free(poly->components);
    poly->components = NULL;
    poly->n_component = 0;
}

void alloc_plaintext(CPlaintext* pt, int level, int n) {
    pt->level = level;
    alloc_poly(&pt->poly, level + 1, n);
}

void alloc_ciphertext(CCiphertext* ct, int degree, int level, int n) {
    ct->degree = degree;
    ct->level = level;
    ct->polys = (CPolynomial*)malloc((size_t)(degree + 1) * sizeof(CPolynomial));
    for (int i = 0; i <= degree; i++) alloc_poly(&ct->polys[i], level + 1, n);
}

void alloc_relin_key(CRelinKey* rlk, int n_public_key, int level, int n) {
    rlk->n_public_key = n_public_key;
    rlk->public_keys = (CCiphertext*)malloc((size_t)n_public_key * sizeof(CCiphertext));
    for (int i = 0; i < n_public_key; i++) alloc_ciphertext(&rlk->public_keys[i], 2, level, n);
}

void set_galois_key_steps(CGaloisKey* glk, uint64_t* galois_elements, int n_galois_elements) {
    glk->n_key_switch_key = n_galois_elements;
    glk->galois_elements = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)n_galois_elements);
    for (int i = 0; i < n_galois_elements; i++) glk->galois_elements[i] = galois_elements[i];
}

void free_plaintext(CPlaintext* pt) { free_poly(&pt->poly); }

void free_ciphertext(CCiphertext* ct) {
    for (int i = 0; i <= ct->degree; i++) free_poly(&ct->polys[i]);
    free(ct->polys);
    ct->polys = NULL;
}

void free_relin_key(CRelinKey* rlk) {
    for (int i = 0; i < rlk->n_public_key; i++) free_ciphertext(&rlk->public_keys[i]);
    free(rlk->public_keys);
    rlk->public_keys = NULL;
}

void free_galois_key(CGaloisKey* gk) {
    for (int i = 0; i < gk->n_key_switch_key; i++) free_relin_key(&gk->key_switch_keys[i]);
    free(gk->galois_elements);
    free(gk->key_switch_keys);
    gk->galois_elements = NULL;
    gk->key_switch_keys = NULL;
}

}  // extern "C"
