// Host-only number theory used when a context builds its device tables: primality, primitive
// roots (Lattigo picks psi = g^((q-1)/2N) with g the smallest generator of Z_q^*), NTT-prime
// search for the BFV auxiliary basis, and a small unsigned bigint for products of moduli.
#pragma once
#include <algorithm>
#include <stdexcept>
#include <vector>
#include "lsb_math.h"

namespace lsb {
namespace host {

inline u64 mulmod(u64 a, u64 b, u64 q) { return (u64)((u128)a * b % q); }
inline u64 powmod(u64 b, u64 e, u64 q) {
    u64 r = 1 % q;
    b %= q;
    for (; e; e >>= 1) {
        if (e & 1) r = mulmod(r, b, q);
        b = mulmod(b, b, q);
    }
    return r;
}
inline u64 invmod(u64 a, u64 q) { return powmod(a % q, q - 2, q); }  // q prime
inline u64 shoup(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
inline Tw make_tw(u64 w, u64 q) { return Tw{w, shoup(w, q)}; }
inline BarrettQ make_barrett(u64 q) {
    u128 r = ~(u128)0 / q;  // floor((2^128-1)/q) == floor(2^128/q) for odd q > 1
    u32 sh = 0;
    while (sh < 63 && (q >> (sh + 1))) sh++;  // bitlen(q) - 1
    if (sh == 0) sh = 1;                      // q < 2 never occurs; keeps the shifts in range
    // q == 2^sh (a power-of-two plaintext modulus) would need mu = 2^64: use 2^64 - 1, the estimate is then low by
    // at most one more unit, still within the two conditional subtractions of reduce_wide
    const u128 mu_full = ((u128)1 << (sh + 64)) / q;
    const u64 mu = (mu_full >> 64) ? ~0ull : (u64)mu_full;
    return BarrettQ{q, (u64)r, (u64)(r >> 64), mu, ~0ull / q, sh, 0};
}

inline bool miller_rabin(u64 n) {
    if (n < 2) return false;
    static const u64 bases[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (u64 p : bases)
        if (n % p == 0) return n == p;
    u64 d = n - 1;
    int s = 0;
    while (!(d & 1)) d >>= 1, s++;
    for (u64 a : bases) {
        u64 x = powmod(a, d, n);
        if (x == 1 || x == n - 1) continue;
        bool witness = true;
        for (int i = 1; i < s && witness; i++) {
            x = mulmod(x, x, n);
            if (x == n - 1) witness = false;
        }
        if (witness) return false;
    }
    return true;
}

inline u64 gcd(u64 a, u64 b) {
    while (b) {
        u64 t = a % b;
        a = b;
        b = t;
    }
    return a;
}
inline u64 rho(u64 n) {
    if (!(n & 1)) return 2;
    for (u64 c = 1;; c++) {
        u64 x = 2, y = 2, d = 1;
        while (d == 1) {
            x = (u64)(((u128)x * x + c) % n);
            y = (u64)(((u128)y * y + c) % n);
            y = (u64)(((u128)y * y + c) % n);
            d = gcd(x > y ? x - y : y - x, n);
        }
        if (d != n) return d;
    }
}
inline void factor_into(u64 n, std::vector<u64>& out) {
    if (n == 1) return;
    if (miller_rabin(n)) {
        out.push_back(n);
        return;
    }
    u64 d = rho(n);
    factor_into(d, out);
    factor_into(n / d, out);
}
inline std::vector<u64> prime_divisors(u64 n) {
    std::vector<u64> f;
    for (u64 p : {2ull, 3ull, 5ull, 7ull, 11ull, 13ull})
        if (n % p == 0) {
            f.push_back(p);
            while (n % p == 0) n /= p;
        }
    factor_into(n, f);
    std::sort(f.begin(), f.end());
    f.erase(std::unique(f.begin(), f.end()), f.end());
    return f;
}
inline u64 smallest_generator(u64 q) {
    const std::vector<u64> f = prime_divisors(q - 1);
    for (u64 g = 2;; g++) {
        bool ok = true;
        for (u64 p : f)
            if (powmod(g, (q - 1) / p, q) == 1) {
                ok = false;
                break;
            }
        if (ok) return g;
    }
}
// primes = 1 mod m walking down from 2^bits + 1 (Lattigo ring.GenerateNTTPrimesP)
inline std::vector<u64> ntt_primes_below(int bits, u64 m, int count) {
    std::vector<u64> out;
    u64 x = (1ull << bits) + 1;
    while ((int)out.size() < count) {
        if (x <= m) throw std::runtime_error("lsb: ran out of NTT prime candidates");
        x -= m;
        if (miller_rabin(x)) out.push_back(x);
    }
    return out;
}
inline u64 bitrev(u64 x, int bits) {
    u64 r = 0;
    for (int i = 0; i < bits; i++) r = (r << 1) | ((x >> i) & 1);
    return r;
}

struct BigU {
    std::vector<u64> w{1, 0};
    explicit BigU(u64 v = 0) : w(1, v) {}
    void mul(u64 m) {
        u64 c = 0;
        for (auto& x : w) {
            u128 t = (u128)x * m + c;
            x = (u64)t;
            c = (u64)(t >> 64);
        }
        if (c) w.push_back(c);
    }
    void shr1() {
        for (size_t i = 0; i < w.size(); i++) w[i] = (w[i] >> 1) | (i + 1 < w.size() ? w[i + 1] << 63 : 0);
        while (w.size() > 1 && !w.back()) w.pop_back();
    }
    u64 mod(u64 m) const {
        u64 r = 0;
        for (size_t i = w.size(); i-- > 0;) r = (u64)((((u128)r << 64) | w[i]) % m);
        return r;
    }
    int bits() const {
        size_t i = w.size() - 1;
        return w[i] ? (int)(i * 64) + 64 - __builtin_clzll(w[i]) : 0;
    }
};

}  // namespace host
}  // namespace lsb
