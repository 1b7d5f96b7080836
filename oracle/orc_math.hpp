// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked, imported or executed by
// the product path (lattisense_b200/); only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs use it, and only as the checker / CPU baseline.
//
// PARITY UNPINNED: the reference's arithmetic lives in the absent submodule
// github.com/cipherflow-fhe/lattigo @ d68e5122 (a fork of tuneinsight/lattigo v4); the reference
// holds no golden vectors or KATs for this path.  This file restates the published
// Lattigo v4 algorithms (ring/primes.go, ring/ntt.go, ring/modular_reduction.go) with plain
// unsigned __int128 arithmetic.
#pragma once
#include <cstdint>
#include <vector>
#include <stdexcept>
#include <algorithm>

namespace orc {

typedef uint64_t u64;
typedef unsigned __int128 u128;

inline u64 mulmod(u64 a, u64 b, u64 q) { return (u64)((u128)a * b % q); }
inline u64 addmod(u64 a, u64 b, u64 q) { u64 s = a + b; return (s >= q || s < a) ? s - q : s; }
inline u64 submod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }
inline u64 negmod(u64 a, u64 q) { return a ? q - a : 0; }

inline u64 powmod(u64 b, u64 e, u64 q) {
    u64 r = 1 % q;
    b %= q;
    while (e) {
        if (e & 1) r = mulmod(r, b, q);
        b = mulmod(b, b, q);
        e >>= 1;
    }
    return r;
}

// q prime
inline u64 invmod(u64 a, u64 q) { return powmod(a % q, q - 2, q); }

inline bool is_prime(u64 n) {
    if (n < 2) return false;
    static const u64 small[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (u64 p : small) {
        if (n % p == 0) return n == p;
    }
    u64 d = n - 1;
    int s = 0;
    while (!(d & 1)) { d >>= 1; s++; }
    for (u64 a : small) {  // deterministic for n < 2^64
        u64 x = powmod(a, d, n);
        if (x == 1 || x == n - 1) continue;
        bool comp = true;
        for (int i = 1; i < s; i++) {
            x = mulmod(x, x, n);
            if (x == n - 1) { comp = false; break; }
        }
        if (comp) return false;
    }
    return true;
}

inline u64 gcd64(u64 a, u64 b) { while (b) { u64 t = a % b; a = b; b = t; } return a; }

inline u64 pollard_rho(u64 n) {
    if (!(n & 1)) return 2;
    for (u64 c = 1;; c++) {
        u64 x = 2, y = 2, d = 1;
        auto f = [&](u64 v) { return (u64)(((u128)v * v + c) % n); };
        while (d == 1) {
            x = f(x);
            y = f(f(y));
            d = gcd64(x > y ? x - y : y - x, n);
        }
        if (d != n) return d;
    }
}

inline void factor_rec(u64 n, std::vector<u64>& out) {
    if (n == 1) return;
    if (is_prime(n)) { out.push_back(n); return; }
    u64 d = pollard_rho(n);
    factor_rec(d, out);
    factor_rec(n / d, out);
}

inline std::vector<u64> distinct_prime_factors(u64 n) {
    std::vector<u64> f;
    for (u64 p : {2ull, 3ull, 5ull, 7ull, 11ull, 13ull}) {
        if (n % p == 0) { f.push_back(p); while (n % p == 0) n /= p; }
    }
    factor_rec(n, f);
    std::sort(f.begin(), f.end());
    f.erase(std::unique(f.begin(), f.end()), f.end());
    return f;
}

// Lattigo v4 ring.PrimitiveRoot: smallest generator g >= 2 of Z_q^*.
inline u64 primitive_root(u64 q) {
    std::vector<u64> fac = distinct_prime_factors(q - 1);
    for (u64 g = 2;; g++) {
        bool ok = true;
        for (u64 f : fac) {
            if (powmod(g, (q - 1) / f, q) == 1) { ok = false; break; }
        }
        if (ok) return g;
    }
}

// Lattigo v4 ring.GenerateNTTPrimesP(logP, NthRoot, n): primes = 1 mod NthRoot, walking down from 2^logP + 1.
inline std::vector<u64> generate_ntt_primes_p(int logP, u64 nth_root, int n) {
    std::vector<u64> primes;
    u64 x = (1ull << logP) + 1;
    while ((int)primes.size() < n) {
        if (x <= nth_root) throw std::runtime_error("generate_ntt_primes_p: ran out of candidates");
        x -= nth_root;
        if (is_prime(x)) primes.push_back(x);
    }
    return primes;
}

inline u64 bitrev(u64 x, int bits) {
    u64 r = 0;
    for (int i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}

// Tiny unsigned bigint (little-endian 64-bit words) — just enough for Q products, Q/2 and residues.
struct BigUInt {
    std::vector<u64> w;
    BigUInt(u64 v = 0) : w(1, v) {}
    void mul(u64 m) {
        u64 carry = 0;
        for (auto& x : w) { u128 t = (u128)x * m + carry; x = (u64)t; carry = (u64)(t >> 64); }
        if (carry) w.push_back(carry);
    }
    void shr1() {
        for (size_t i = 0; i < w.size(); i++) {
            w[i] >>= 1;
            if (i + 1 < w.size()) w[i] |= w[i + 1] << 63;
        }
        while (w.size() > 1 && w.back() == 0) w.pop_back();
    }
    u64 mod(u64 m) const {
        u64 r = 0;
        for (size_t i = w.size(); i-- > 0;) r = (u64)((((u128)r << 64) | w[i]) % m);
        return r;
    }
    int bitlen() const {
        size_t i = w.size() - 1;
        if (w[i] == 0) return 0;
        return (int)(i * 64) + (64 - __builtin_clzll(w[i]));
    }
};

}  // namespace orc
