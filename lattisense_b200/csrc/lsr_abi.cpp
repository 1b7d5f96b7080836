// Allocation helpers for the ABI structs (same ownership rules as the reference's abi/c_structs.c:23-98:
// one malloc per limb, freed limb by limb).  The Lattigo import/export shims of that file are not part of
// this library: they belong to the caller that links liblattigo.
#include <stdlib.h>

#include "lattisense_gpu_runner.h"

extern "C" {

void alloc_component(CComponent* component, int n) {
    component->n = n;
    component->data = (uint64_t*)malloc((size_t)n * sizeof(uint64_t));
}

static void alloc_poly(CPolynomial* poly, int n_component, int n) {
    poly->n_component = n_component;
    poly->components = (CComponent*)malloc((size_t)n_component * sizeof(CComponent));
    for (int i = 0; i < n_component; i++) alloc_component(&poly->components[i], n);
}
static void free_poly(CPolynomial* poly) {
    for (int i = 0; i < poly->n_component; i++) free(poly->components[i].data);
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
