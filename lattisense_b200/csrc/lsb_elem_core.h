// Per-coefficient bodies of the basis-conversion kernels, LSB_HD so tests/emul can run them on the
// CPU against the oracle.  Algorithm: Lattigo v4 ring/basis_extension.go (modUpExact /
// reconstructRNS with a float64 overflow count), restated in oracle/orc_core.hpp:139-202.
#pragma once
#include "lsb_kernels.h"
#include "lsb_ntt_arith.h"

namespace lsb {

// x[i]: source residues of one coefficient.  Writes job.nt outputs through `store(t, value)`;
// `other(t)` supplies the extra operand of BEX_OTHER_SUB.
template <class Other, class Store>
LSB_HD void bex_coeff(const BexJob& job, const ModConst* mods, const u64* x, Other&& other, Store&& store) {
    if (job.centered) {
        const u64 s = mods[job.src_mod[0]].q;
        u64 c = x[0];
        const bool neg = c >= (s >> 1);
        if (neg) c = s - c;
        for (int t = 0; t < job.nt; t++) {
            const ModConst& m = mods[job.dst_mod[t]];
            u64 r = barrett64(c, m.bq);
            store(t, neg ? negmod(r, m.q) : r);
        }
        return;
    }
    if (job.f64_ok) {
        // every modulus below 2^47: the same CRT on the FP64 pipe (the 64-bit integer multiplies of the path below
        // are what bounds this kernel on B200).  y_i is brought to its canonical value first, so the float64
        // overflow estimate sums exactly the terms Lattigo sums.
        double yd[kBexMaxSrc];
        double vf = 0.0;
        for (int i = 0; i < job.ns; i++) {
            const ModConst& ms = mods[job.src_mod[i]];
            yd[i] = f64_canon_d(f64_mulmod(f64_from_u64(x[i]), job.qhat_inv_d[i], ms.qd), ms.qd, ms.qinvd);
            vf += yd[i] / job.s_f64[i];
        }
        const int v = (int)(u64)vf;
        for (int t = 0; t < job.nt; t++) {
            const ModConst& m = mods[job.dst_mod[t]];
            double acc = 0.0;  // |acc| <= 0.75 * ns * p < 2^51
            for (int i = 0; i < job.ns; i++) acc = d_add(acc, f64_mulmod(yd[i], job.qhat_mod_d[t][i], m.qd));
            acc = d_add(acc, -f64_from_u64(job.vS[t][v]));
            if (job.mode == BEX_OTHER_SUB) {
                const double e = f64_canon_d(acc, m.qd, m.qinvd);
                const double d = d_add(f64_from_u64(other(t)), -e);
                acc = d_add(f64_mulmod(d, job.mulc_d[t], m.qd), f64_from_u64(job.addc[t]));
            } else if (job.mode == BEX_SUB_MUL) {
                const double e = f64_canon_d(acc, m.qd, m.qinvd);
                acc = f64_mulmod(d_add(e, -f64_from_u64(job.addc[t])), job.mulc_d[t], m.qd);
            }
            store(t, f64_canon_u(acc, m.qd, m.qinvd));
        }
        return;
    }
    u64 y[kBexMaxSrc];
    // Overflow count = floor of Lattigo's float64 sum v = (..(y_0/s_0) + y_1/s_1 ..), divisions in limb order.  A double
    // division is ~25 instructions, so the sum is first formed with multiplications by 1/s_i: each term is below 1 and the two
    // sums differ by less than ns * 2^-51, hence have the same floor unless the fast one lies within 2^-40 of an integer;
    // only then (probability ~2^-39 per coefficient) is the exact sequence run (the same scheme as k_bex_f64).
    double yd[kBexMaxSrc];
    double vm = 0.0;
    for (int i = 0; i < job.ns; i++) {
        y[i] = mul_shoup(x[i], job.qhat_inv[i], mods[job.src_mod[i]].q);
        yd[i] = (double)y[i];
        vm = d_fma(yd[i], job.rs_f64[i], vm);
    }
    int v = (int)vm;
    {
        const double eps = 9.094947017729282e-13;  // 2^-40
        const double fr = vm - (double)v;
        if (fr < eps || fr > 1.0 - eps) {
            double vf = 0.0;
            for (int i = 0; i < job.ns; i++) vf += yd[i] / job.s_f64[i];
            v = (int)(u64)vf;
        }
    }
    double yl[kBexMaxSrc], yh[kBexMaxSrc];
    if (job.n_f64 > 0) {
        for (int i = 0; i < job.ns; i++) {
            if (job.src_split) {
                yl[i] = f64_from_u64(y[i] & 0xFFFFFFFFull);
                yh[i] = f64_from_u64(y[i] >> 32);
            } else {
                yl[i] = f64_from_u64(y[i]);
            }
        }
    }
    for (int t = 0; t < job.nt; t++) {
        const ModConst& m = mods[job.dst_mod[t]];
        if (job.t_f64[t]) {
            // this target on the FP64 pipe (mixed-pipe CRT, lsb_tables.h bex_choose_mixed): the same sum of y_i * (S/s_i mod p),
            // each product reduced to [-0.75 p, 0.75 p], the result brought to the canonical residue -- identical limbs
            double acc = 0.0;
            for (int i = 0; i < job.ns; i++) {
                acc = d_add(acc, f64_mulmod(yl[i], job.qhat_mod_d[t][i], m.qd));
                if (job.src_split) acc = d_add(acc, f64_mulmod(yh[i], job.qhat_mod_hi_d[t][i], m.qd));
            }
            acc = d_add(acc, -f64_from_u64(job.vS[t][v]));
            if (job.mode == BEX_OTHER_SUB) {
                const double e = f64_canon_d(acc, m.qd, m.qinvd);
                const double d = d_add(f64_from_u64(other(t)), -e);
                acc = d_add(f64_mulmod(d, job.mulc_d[t], m.qd), f64_from_u64(job.addc[t]));
            } else if (job.mode == BEX_SUB_MUL) {
                const double e = f64_canon_d(acc, m.qd, m.qinvd);
                acc = f64_mulmod(d_add(e, -f64_from_u64(job.addc[t])), job.mulc_d[t], m.qd);
            }
            store(t, f64_canon_u(acc, m.qd, m.qinvd));
            continue;
        }
        Acc128 acc{0, 0};
        for (int i = 0; i < job.ns; i++) acc_mad(acc, y[i], job.qhat_mod[t][i]);
        u64 e = job.wide_ok ? reduce_wide(acc.lo, acc.hi, m.bq) : barrett128(acc.lo, acc.hi, m.bq);
        e = submod(e, job.vS[t][v], m.q);
        if (job.mode == BEX_OTHER_SUB) {
            u64 d = submod(other(t), e, m.q);
            e = addmod(mul_shoup(d, job.mulc[t], m.q), job.addc[t], m.q);
        } else if (job.mode == BEX_SUB_MUL) {
            e = mul_shoup(submod(e, job.addc[t], m.q), job.mulc[t], m.q);
        }
        store(t, e);
    }
}

}  // namespace lsb
