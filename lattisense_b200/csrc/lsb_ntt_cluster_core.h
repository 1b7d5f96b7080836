// Shared-memory round of the cluster NTT (LSB_HD so that tests/emul can run it on the CPU).
#pragma once
#include "lsb_ntt_arith.h"
#include "lsb_ntt_core.h"

namespace lsb {

// radix (log2) of the contiguous last local round of the cluster transform for ring degree 2^logN
LSB_HD constexpr int cluster_last_lr(int logN) {
    const Plan p = make_plan(logN - 4, 0);
    return p.rounds[p.nrounds - 1].lr;
}

// one local round on sub-tile `smt` (BS words, swizzled): LR stages starting at local stage v.
// RAW_SRC (first inverse round): canonical residues are read straight from global memory `gsrc` (block base);
// CANON (last forward round): canonical residues are written to global memory `gdst` if given (measured
// slower than staging through shared memory on B200: half-sector stores), else to shared memory.  Both only
// occur in the contiguous round (logGC == 0), where a thread owns 2^LR consecutive words.
// S1 >= 0: tail-SoA twiddle layout (lsb_ntt_arith.h), only valid for the contiguous round whose first global stage is S1.
template <class P, int LR, bool INV, bool RAW_SRC, bool CANON, int S1 = -1>
LSB_HD void local_round(u64* smt, const typename P::TW* __restrict__ tw, int g, int logGC, int v, u32 twbase,
                        const ModCtx& m, const u64* __restrict__ gsrc = nullptr, u64* __restrict__ gdst = nullptr) {
    constexpr int R = 1 << LR;
    const int hi = g >> logGC;
    const int u0 = (hi << (LR + logGC)) + (g & ((1 << logGC) - 1));
    typename P::V x[R];
    if (logGC == 0) {
#pragma unroll
        for (int k = 0; k < R; k += 2) {
            const W2 pr = RAW_SRC ? *reinterpret_cast<const W2*>(&gsrc[u0 + k]) : *reinterpret_cast<const W2*>(&smt[swz(u0 + k)]);
            x[k] = RAW_SRC ? P::from_raw(pr.a) : P::from_bits(pr.a);
            x[k + 1] = RAW_SRC ? P::from_raw(pr.b) : P::from_bits(pr.b);
        }
    } else {
#pragma unroll
        for (int k = 0; k < R; k++) {
            const u64 b = smt[swz(u0 + (k << logGC))];
            x[k] = RAW_SRC ? P::from_raw(b) : P::from_bits(b);
        }
    }
    const u32 tbase = (twbase << v) + (u32)hi;
    if (!INV) fwd_stages<P, LR, S1>(x, tw, tbase, m);
    else inv_stages<P, LR, false, S1>(x, tw, tbase, m, typename P::TW{}, typename P::TW{});
    if (logGC == 0) {
#pragma unroll
        for (int k = 0; k < R; k += 2) {
            W2 pr;
            pr.a = CANON ? P::fwd_canonical(x[k], m) : P::to_bits(x[k]);
            pr.b = CANON ? P::fwd_canonical(x[k + 1], m) : P::to_bits(x[k + 1]);
            if (CANON && gdst) *reinterpret_cast<W2*>(&gdst[u0 + k]) = pr;
            else *reinterpret_cast<W2*>(&smt[swz(u0 + k)]) = pr;
        }
    } else {
#pragma unroll
        for (int k = 0; k < R; k++) smt[swz(u0 + (k << logGC))] = CANON ? P::fwd_canonical(x[k], m) : P::to_bits(x[k]);
    }
}

}  // namespace lsb
