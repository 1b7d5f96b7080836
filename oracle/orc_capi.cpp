// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.hpp header).  C ABI over the CPU restatement so
// that tests/ (ctypes) and bench.py's cpu_baseline leg can call it.  Flat u64 arrays, explicit shapes:
//   ct  [deg+1][level+1][N]      evk [beta][2][key_level+1+k][N]
#include "orc_keys.hpp"

using namespace orc;

extern "C" {

void* orc_create(int scheme, int logN, const u64* q, int nq, const u64* p, int np, u64 t) {
    try {
        return make_context(scheme, logN, q, nq, p, np, t).release();
    } catch (const std::exception&) {
        return nullptr;
    }
}
void orc_destroy(void* h) { delete (Context*)h; }

int orc_num_moduli(void* h) { return (int)((Context*)h)->mods.size(); }
int orc_num_qmul(void* h) { return ((Context*)h)->nqm; }
int orc_level_qmul(void* h, int level) { return ((Context*)h)->level_qmul.at(level); }
u64 orc_modulus(void* h, int midx) { return ((Context*)h)->mods.at(midx).q; }
u64 orc_psi(void* h, int midx) { return ((Context*)h)->mods.at(midx).psi; }

void orc_ntt(void* h, int midx, u64* a, int inverse) {
    Context& c = *(Context*)h;
    if (inverse) ntt_inverse(c.mods.at(midx), c.N, a);
    else ntt_forward(c.mods.at(midx), c.N, a);
}

void orc_permute_ntt_index(int logN, u64 gal, uint32_t* out) {
    auto v = permute_ntt_index(logN, gal);
    std::memcpy(out, v.data(), v.size() * sizeof(uint32_t));
}

// generic Lattigo-style basis extension between arbitrary modulus-index sets; src [ns][N], dst [nt][N]
void orc_basis_extend(void* h, const int* sidx, int ns, const u64* src, const int* tidx, int nt, u64* dst) {
    Context& c = *(Context*)h;
    std::vector<const u64*> s(ns);
    std::vector<u64*> d(nt);
    for (int i = 0; i < ns; i++) s[i] = src + (size_t)i * c.N;
    for (int j = 0; j < nt; j++) d[j] = dst + (size_t)j * c.N;
    basis_extend(c, sidx, ns, s.data(), tidx, nt, d.data());
}

void orc_keyswitch_ntt(void* h, int level, const u64* c_ntt, const u64* evk, int key_level, u64* p0, u64* p1) {
    keyswitch_ntt(*(Context*)h, level, c_ntt, evk, key_level, p0, p1);
}

void orc_add(void* h, int level, int npoly, const u64* a, const u64* b, u64* o) { ct_add(*(Context*)h, level, npoly, a, b, o); }
void orc_sub(void* h, int level, int npoly, const u64* a, const u64* b, u64* o) { ct_sub(*(Context*)h, level, npoly, a, b, o); }
void orc_neg(void* h, int level, int npoly, const u64* a, u64* o) { ct_neg(*(Context*)h, level, npoly, a, o); }

void orc_mult(void* h, int level, const u64* a, const u64* b, u64* out3) {
    Context& c = *(Context*)h;
    if (c.scheme == SCHEME_CKKS) ckks_mult(c, level, a, b, out3);
    else bfv_mult(c, level, a, b, out3);
}
void orc_relin(void* h, int level, const u64* ct3, const u64* rlk, int key_level, u64* out2) {
    Context& c = *(Context*)h;
    if (c.scheme == SCHEME_CKKS) ckks_relin(c, level, ct3, rlk, key_level, out2);
    else bfv_relin(c, level, ct3, rlk, key_level, out2);
}
void orc_rescale(void* h, int level, int npoly, const u64* in, u64* out) {
    Context& c = *(Context*)h;
    if (c.scheme == SCHEME_CKKS) ckks_rescale(c, level, npoly, in, out);
    else bfv_rescale(c, level, npoly, in, out);
}
void orc_drop_level(void* h, int level, int npoly, const u64* in, u64* out) { drop_level(*(Context*)h, level, npoly, in, out); }
void orc_rotate(void* h, int level, const u64* ct, const u64* glk, int key_level, u64 gal, u64* out) {
    Context& c = *(Context*)h;
    if (c.scheme == SCHEME_CKKS) ckks_rotate(c, level, ct, glk, key_level, gal, out);
    else bfv_rotate(c, level, ct, glk, key_level, gal, out);
}

// ---- test-only client-side crypto ----
void orc_gen_secret(void* h, u64 seed, int8_t* s) { gen_secret(*(Context*)h, seed, s); }
void orc_encrypt_sk(void* h, const int8_t* s, u64 seed, int level, const u64* pt, u64* ct) {
    encrypt_sk(*(Context*)h, s, seed, level, pt, ct);
}
void orc_decrypt(void* h, const int8_t* s, int level, int degree, const u64* ct, u64* pt) {
    decrypt(*(Context*)h, s, level, degree, ct, pt);
}
void orc_gen_rlk(void* h, const int8_t* s, u64 seed, int key_level, u64* evk) { gen_rlk(*(Context*)h, s, seed, key_level, evk); }
void orc_gen_glk(void* h, const int8_t* s, u64 seed, int key_level, u64 gal, u64* evk) {
    gen_glk(*(Context*)h, s, seed, key_level, gal, evk);
}
void orc_bfv_scale_up(void* h, int level, const u64* m, u64* pt) { bfv_scale_up(*(Context*)h, level, m, pt); }

}  // extern "C"
