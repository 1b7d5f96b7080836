// Negacyclic NTT tile engine (shared-memory staged radix-8/16 rounds).
//
// Ordering follows the reference's ABI domain (Lattigo ring.NTT: Cooley-Tukey, natural in ->
// bit-reversed out; ring.InvNTT: Gentleman-Sande with N^-1 folded; SURVEY.md fact 6,
// plug-in/SEAL/acc/c_struct_import_export.h:39-43).  Stage s (m = 2^s groups, half-gap t = N/2m)
// pairs (j, j+t) and uses twiddle psi_br[m + j/2t].
//
// A "tile" is S*C words held in shared memory: S = 2^LS points of a sub-transform along the
// leading dimension, C = 2^LC independent columns, logical index u = a*C + c.  A "round" runs LR
// consecutive stages on R = 2^LR words held in registers: group g owns
//     u_k = (g / GC) * R * GC + k * GC + (g % GC),   GC = (S >> (v+LR)) * C,  k = 0..R-1
// where v is the first (tile-local) stage of the round.  For the global stage s = s0 + v + r the
// twiddle index is ((twbase << v) + g/GC) << r | (k >> (LR-r)) with twbase = 2^s0 + (tile's
// leading index).  Shared memory is addressed through a 128-byte XOR swizzle so that both the
// strided rounds (GC >= 8) and the contiguous last round (GC = 1, 128-bit accesses) are
// bank-conflict free.
//
// All functions are LSB_HD: tests/emul runs the very same round code on the CPU (thread loops in
// place of a CTA) to validate indexing against the oracle without a GPU.
#pragma once
#include "lsb_math.h"

namespace lsb {

LSB_HD int swz(int u) { return u ^ (((u >> 4) & 7) << 1); }

// Global address of tile word u: ((u >> LC) << gshift) + (u & (C-1)).  Contiguous tiles use LC = 0, gshift = 0.
struct TileGeom {
    int lc;      // log2 columns
    int gshift;  // log2 of the global stride between consecutive leading indices
};
LSB_HD size_t gaddr(const TileGeom& tg, int u) {
    return ((size_t)(u >> tg.lc) << tg.gshift) + (size_t)(u & ((1 << tg.lc) - 1));
}

enum { IO_SMEM = 0, IO_GLOBAL = 1 };
struct __attribute__((aligned(16))) W2 {
    u64 a, b;
};

// One forward round.  SRC/DST select shared memory (swizzled) or global memory (through gaddr).
// FINAL: reduce outputs to [0,q) (last round of the whole transform).
// LAZY: no per-butterfly range correction (q < kLazyLimit); FINAL then reduces with qinv64.
template <int LR, int SRC, int DST, bool FINAL, bool VEC, bool LAZY = false>
LSB_HD void fwd_round(u64* sm, const u64* gsrc, u64* gdst, const TileGeom tg, const Tw* __restrict__ tw, int g,
                      int logGC, int v, u32 twbase, u64 q, u64 qinv64 = 0) {
    constexpr int R = 1 << LR;
    const u64 q2 = q << 1;
    const int hi = g >> logGC;
    const int u0 = (hi << (LR + logGC)) + (g & ((1 << logGC) - 1));
    u64 x[R];
    if (SRC == IO_SMEM && VEC) {  // logGC == 0: R contiguous words, 128-bit accesses
#pragma unroll
        for (int k = 0; k < R; k += 2) {
            const W2 pr = *reinterpret_cast<const W2*>(&sm[swz(u0 + k)]);
            x[k] = pr.a;
            x[k + 1] = pr.b;
        }
    } else {
#pragma unroll
        for (int k = 0; k < R; k++) {
            int u = u0 + (k << logGC);
            x[k] = (SRC == IO_GLOBAL) ? gsrc[gaddr(tg, u)] : sm[swz(u)];
        }
    }
    const u32 tbase = (twbase << v) + (u32)hi;
#pragma unroll
    for (int r = 0; r < LR; r++) {
        const int hk = R >> (r + 1);
#pragma unroll
        for (int kg = 0; kg < (1 << r); kg++) {
            const Tw t = tw[(tbase << r) | (u32)kg];
#pragma unroll
            for (int i = 0; i < hk; i++) {
                if (LAZY) ct_bfly_nocorr(x[kg * 2 * hk + i], x[kg * 2 * hk + i + hk], t, q, q2);
                else ct_bfly(x[kg * 2 * hk + i], x[kg * 2 * hk + i + hk], t, q, q2);
            }
        }
    }
    if (FINAL) {
#pragma unroll
        for (int k = 0; k < R; k++) x[k] = LAZY ? reduce_full(x[k], q, qinv64) : csub(csub(x[k], q2), q);
    }
    if (DST == IO_SMEM && VEC) {
#pragma unroll
        for (int k = 0; k < R; k += 2) {
            W2 pr;
            pr.a = x[k];
            pr.b = x[k + 1];
            *reinterpret_cast<W2*>(&sm[swz(u0 + k)]) = pr;
        }
    } else {
#pragma unroll
        for (int k = 0; k < R; k++) {
            int u = u0 + (k << logGC);
            if (DST == IO_GLOBAL) gdst[gaddr(tg, u)] = x[k];
            else sm[swz(u)] = x[k];
        }
    }
}

// One inverse round (stages executed in reverse order).  FINAL: the round contains global stage 0;
// N^-1 is folded into that stage and outputs are reduced to [0,q).
template <int LR, int SRC, int DST, bool FINAL, bool VEC>
LSB_HD void inv_round(u64* sm, const u64* gsrc, u64* gdst, const TileGeom tg, const Tw* __restrict__ tw, int g,
                      int logGC, int v, u32 twbase, u64 q, Tw ninv, Tw w_ninv) {
    constexpr int R = 1 << LR;
    const u64 q2 = q << 1;
    const int hi = g >> logGC;
    const int u0 = (hi << (LR + logGC)) + (g & ((1 << logGC) - 1));
    u64 x[R];
    if (SRC == IO_SMEM && VEC) {  // logGC == 0: R contiguous words, 128-bit accesses
#pragma unroll
        for (int k = 0; k < R; k += 2) {
            const W2 pr = *reinterpret_cast<const W2*>(&sm[swz(u0 + k)]);
            x[k] = pr.a;
            x[k + 1] = pr.b;
        }
    } else {
#pragma unroll
        for (int k = 0; k < R; k++) {
            int u = u0 + (k << logGC);
            x[k] = (SRC == IO_GLOBAL) ? gsrc[gaddr(tg, u)] : sm[swz(u)];
        }
    }
    const u32 tbase = (twbase << v) + (u32)hi;
#pragma unroll
    for (int r = LR - 1; r >= 0; r--) {
        const int hk = R >> (r + 1);
        if (FINAL && r == 0) {
            // global stage 0: x' = (U+V) * N^-1, y' = (U-V) * psi^-1_br[1] * N^-1
#pragma unroll
            for (int i = 0; i < hk; i++) {
                u64 U = x[i], V = x[i + hk];
                x[i] = mul_shoup(U + V, ninv, q);
                x[i + hk] = mul_shoup(U - V + q2, w_ninv, q);
            }
        } else {
#pragma unroll
            for (int kg = 0; kg < (1 << r); kg++) {
                const Tw t = tw[(tbase << r) | (u32)kg];
#pragma unroll
                for (int i = 0; i < hk; i++) gs_bfly(x[kg * 2 * hk + i], x[kg * 2 * hk + i + hk], t, q, q2);
            }
        }
    }
    if (DST == IO_SMEM && VEC) {
#pragma unroll
        for (int k = 0; k < R; k += 2) {
            W2 pr;
            pr.a = x[k];
            pr.b = x[k + 1];
            *reinterpret_cast<W2*>(&sm[swz(u0 + k)]) = pr;
        }
    } else {
#pragma unroll
        for (int k = 0; k < R; k++) {
            int u = u0 + (k << logGC);
            if (DST == IO_GLOBAL) gdst[gaddr(tg, u)] = x[k];
            else sm[swz(u)] = x[k];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Transform plans.  A plan is a list of rounds (first tile-local stage v, log2 radix lr); GC is
// derived.  Single-pass plans keep a whole limb in shared memory (N <= 2^14: 128 KiB); larger
// rings use two passes (column pass over strided tiles, then row pass over contiguous tiles).
// ---------------------------------------------------------------------------------------------
struct Round {
    int v, lr;
};
struct Plan {
    int ls;  // log2 points of the tile's sub-transform
    int lc;  // log2 columns
    int nrounds;
    Round rounds[5];
};

LSB_HD constexpr Plan make_plan(int ls, int lc) {
    Plan p{};
    p.ls = ls;
    p.lc = lc;
    p.nrounds = 0;
    // ls = 4*n4 + 3*n3 with the radix-16 rounds first, so that the late (small-gap) rounds are
    // radix 8: GC = 8 then GC = 1, the two geometries the XOR swizzle serves without conflicts.
    int n4 = -1;
    for (int c = 0; 4 * c <= ls; c++)
        if ((ls - 4 * c) % 3 == 0) {
            n4 = c;
            break;
        }
    int v = 0, rem = ls;
    if (n4 < 0) {  // ls in {1, 2, 5}: one short round first
        int head = ls % 3;
        p.rounds[p.nrounds].v = v;
        p.rounds[p.nrounds].lr = head;
        p.nrounds++;
        v += head;
        rem -= head;
        n4 = 0;
    }
    for (int i = 0; i < n4; i++) {
        p.rounds[p.nrounds].v = v;
        p.rounds[p.nrounds].lr = 4;
        p.nrounds++;
        v += 4;
        rem -= 4;
    }
    while (rem > 0) {
        p.rounds[p.nrounds].v = v;
        p.rounds[p.nrounds].lr = 3;
        p.nrounds++;
        v += 3;
        rem -= 3;
    }
    return p;
}
// log2 of the contiguous run owned by consecutive groups in round i
LSB_HD constexpr int plan_loggc(const Plan& p, int i) { return p.ls - p.rounds[i].v - p.rounds[i].lr + p.lc; }

}  // namespace lsb
