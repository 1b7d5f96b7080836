// Index helpers of the persistent cluster NTT (lsb_ntt_persist.cu), LSB_HD so that tests/emul checks them on the CPU.
//
// The persistent kernel keeps the twiddles of the 1/16 blocks a CTA owns in SHARED memory for as long as it works on
// limbs of one modulus.  A block's local rounds (stages 4 .. logN-1 of the whole transform) only touch the sub-table
//     plain[((16 + gb) << ls) + e],   ls = local stage 0 .. LS-1,  e < 2^ls,   gb = block index 0..15,
// of the modulus' bit-reversed table, i.e. 2^LS - 1 entries.  They are stored as a stand-alone table of BS = 2^LS
// entries indexed as if the block were a whole transform (twbase = 1): entry (1 << ls) + e, with the stages of the
// contiguous last round (local stages s1l = LS - lrl ..) in the tail-SoA order of lsb_ntt_arith.h (tw_index<S1> with
// S1 = s1l), so that consecutive threads read consecutive 16-byte entries from shared memory.
#pragma once
#include "lsb_ntt_cluster_core.h"

namespace lsb {

LSB_HD int persist_log2(u32 v) {
    int l = 0;
    while ((v >> (l + 1)) != 0) l++;
    return l;
}
// local table index (1 <= idx < 2^LS) -> index into the modulus' PLAIN bit-reversed table
LSB_HD u32 persist_tw_src(u32 idx, int gb, int LS, int lrl) {
    const int s1l = LS - lrl;
    int ls;
    u32 e;
    if (idx < ((u32)1 << s1l)) {
        ls = persist_log2(idx);
        e = idx - ((u32)1 << ls);
    } else {
        const u32 top = idx >> s1l, g = idx & (((u32)1 << s1l) - 1);
        const int r = persist_log2(top);
        const u32 kg = top - ((u32)1 << r);
        ls = s1l + r;
        e = (g << r) | kg;
    }
    return (((u32)16 + (u32)gb) << ls) + e;
}

}  // namespace lsb
