// Limb-sharded hybrid key-switch (third mode; BASELINE north_star: "RNS limbs are sharded with an NCCL
// all-gather over NVLink for the ModUp basis extension").
//
// World of G ranks, one GPU each.  Rank r owns the Q limbs {i : i mod G == r} and the P limbs {j : j mod G == r} of
// every polynomial AND of every evaluation key (key memory per GPU drops by G: 150 MiB -> 19 MiB per key at the
// bootstrap parameters).  Limb-wise work (NTT, inner product with the key, finish) needs no communication; the two
// basis conversions need all source limbs of a coefficient, which gives exactly two exchange steps:
//
//   phase 1  begin : INTT of the owned Q limbs of c                       -> coeff_own [cq][N]
//            -- all-gather coeff_own over the ranks                        -> coeff_all [G][cq][N]
//   phase 2  inner : ModUp from coeff_all to the OWNED targets only, NTT, inner product with the owned key limbs,
//                    INTT of the owned P limbs of the accumulators         -> acc_own [2][n_own][N], accP_own [2][cp][N]
//            -- all-gather accP_own                                        -> accP_all [G][2][cp][N]
//   phase 3  end   : ModDown from accP_all to the owned Q limbs, NTT with the fused finish (+ c0/c1 limbs)
//                                                                          -> out_own [2][nq_own][N]
//
// cq = ceil(l/G), cp = ceil(k/G) (ranks that own fewer limbs leave the padding rows untouched).  The collectives are
// issued by the caller (lattisense_b200/dist.py: torch.distributed all_gather_into_tensor over NCCL on the same
// stream); every kernel is the single-GPU one driven by restricted row maps, so each rank's limbs are bit-identical
// to the corresponding limbs of the single-GPU key-switch (tests/test_shard_gpu.py emulates G ranks on one GPU).
#include <algorithm>
#include "lsb_ops.h"
#include "lsb_shard.h"
#include "lsb_tables.h"

namespace lsb {

using namespace host;

ShardPlan::ShardPlan(Context& c, int level_, int world_, int rank_) : level(level_), world(world_), rank(rank_) {
    LSB_CHECK(world >= 1 && rank >= 0 && rank < world, "lsb: bad shard rank / world");
    LSB_CHECK(level >= 0 && level < c.nq, "lsb: level outside the modulus chain");
    l = level + 1;
    k = c.np;
    beta = c.beta(level);
    cq = (l + world - 1) / world;
    cp = (k + world - 1) / world;
    for (int i = rank; i < l; i += world) own_q.push_back(i);
    for (int j = rank; j < k; j += world) own_p.push_back(j);
    nq_own = (int)own_q.size();
    np_own = (int)own_p.size();
    n_own = nq_own + np_own;
    LSB_CHECK(n_own <= 64 && world * std::max(cq, 2 * cp) < 256 && beta * std::max(n_own, 1) < 256,
              "lsb: shard layout exceeds the slot range");
    const int alpha = k;

    // ---- ModUp jobs: digit d, sources in the gathered layout, targets = owned limbs outside the digit ----
    std::vector<BexJob> up;
    ext_rows.n = 0;
    for (int d = 0; d < beta; d++) {
        const int st = d * alpha, ed = std::min(st + alpha, l);
        std::vector<int> src, dst, slots;
        for (int i = st; i < ed; i++) src.push_back(i);
        for (int u = 0; u < nq_own; u++)
            if (own_q[u] < st || own_q[u] >= ed) {
                dst.push_back(own_q[u]);
                slots.push_back(d * n_own + u);
            }
        for (int u = 0; u < np_own; u++) {
            dst.push_back(c.idxP(own_p[u]));
            slots.push_back(d * n_own + nq_own + u);
        }
        if (dst.empty()) continue;
        BexJob job = make_bex_tables(c.moduli, src, dst);
        job.centered = (ed - st == 1) ? 1 : 0;
        for (int i = 0; i < job.ns; i++) job.src_slot[i] = (uint8_t)(((st + i) % world) * cq + (st + i) / world);
        for (int t = 0; t < job.nt; t++) {
            job.dst_slot[t] = (uint8_t)slots[t];
            ext_rows.in_slot[ext_rows.n] = ext_rows.out_slot[ext_rows.n] = (uint8_t)slots[t];
            ext_rows.mod[ext_rows.n] = (uint8_t)dst[t];
            ext_rows.n++;
        }
        up.push_back(job);
    }
    n_up = (int)up.size();
    for (const BexJob& j : up) up_meta.push_back(BexMeta{j.ns, j.nt, j.mode, j.centered, j.f64_ok});
    up_jobs = n_up ? (const BexJob*)c.upload_raw(up.data(), up.size() * sizeof(BexJob)) : nullptr;

    // ---- ModDown jobs: all k P limbs (gathered layout [G][2][cp]) -> owned Q limbs ----
    std::vector<BexJob> down;
    if (nq_own > 0) {
        std::vector<int> src, dst;
        for (int j = 0; j < k; j++) src.push_back(c.idxP(j));
        for (int u = 0; u < nq_own; u++) dst.push_back(own_q[u]);
        for (int p = 0; p < 2; p++) {
            BexJob job = make_bex_tables(c.moduli, src, dst);
            for (int j = 0; j < k; j++) job.src_slot[j] = (uint8_t)(((j % world) * 2 + p) * cp + j / world);
            for (int u = 0; u < nq_own; u++) job.dst_slot[u] = (uint8_t)(p * nq_own + u);
            down.push_back(job);
        }
    }
    n_down = (int)down.size();
    for (const BexJob& j : down) down_meta.push_back(BexMeta{j.ns, j.nt, j.mode, j.centered, j.f64_ok});
    down_jobs = n_down ? (const BexJob*)c.upload_raw(down.data(), down.size() * sizeof(BexJob)) : nullptr;

    // ---- finish maps (with / without the c0, c1 addend) ----
    for (int add = 0; add < 2; add++) {
        FinishMap& fm = fin_host[add];
        std::memset(&fm, 0, sizeof(fm));
        fm.n = 2 * nq_own;
        for (int p = 0; p < 2; p++)
            for (int u = 0; u < nq_own; u++) {
                const int r = p * nq_own + u;
                const u64 qi = c.moduli[own_q[u]];
                u64 pm = 1;
                for (int j = 0; j < k; j++) pm = mulmod(pm, c.moduli[c.idxP(j)] % qi, qi);
                fm.a_slot[r] = (uint8_t)(p * n_own + u);
                fm.b_slot[r] = (uint8_t)r;
                fm.add_slot[r] = add ? (uint8_t)r : (uint8_t)0xFF;
                fm.out_slot[r] = (uint8_t)r;
                fm.mod[r] = (uint8_t)own_q[u];
                fm.c[r] = make_tw(invmod(pm, qi), qi);
            }
        fin_dev[add] = fm.n ? (const FinishMap*)c.upload_raw(&fm, sizeof(fm)) : nullptr;
    }
}

const ShardPlan& Context::shard_plan(int level, int world, int rank) {
    std::lock_guard<std::mutex> lk(shard_mu_);
    const u64 key = ((u64)level << 32) | ((u64)world << 16) | (u64)rank;
    auto it = shard_plans_.find(key);
    if (it == shard_plans_.end()) it = shard_plans_.emplace(key, std::unique_ptr<ShardPlan>(new ShardPlan(*this, level, world, rank))).first;
    return *it->second;
}

// phase 1: coeff_own[u] = INTT(c_own_ntt[u]) for the owned Q limbs.  With `peers` (the coeff_all buffer of every
// rank, peer-mapped) the transform stores its outputs straight into slot rank*cq + u of each of them: the first
// all-gather happens inside the INTT kernel, over NVLink, and coeff_own is not written.
void op_ks_shard_begin(OpEnv& e, const ShardPlan& sp, BatchPtr c_own_ntt, BatchPtr coeff_own, int batch, u64* const* peers) {
    LSB_CUDA(cudaSetDevice(e.c.device));
    if (sp.nq_own == 0) return;
    RowMap m;
    m.n = sp.nq_own;
    for (int u = 0; u < sp.nq_own; u++) {
        m.in_slot[u] = (uint8_t)u;
        m.out_slot[u] = (uint8_t)(peers ? sp.rank * sp.cq + u : u);
        m.mod[u] = (uint8_t)sp.own_q[u];
    }
    if (peers) {
        LSB_CHECK(batch == 1, "lsb: the fused all-gather handles one ciphertext per call");
        launch_intt_to_peers(e.c.dt, c_own_ntt, m, peers, sp.world, e.s);
    } else {
        launch_ntt(e.c.dt, c_own_ntt, coeff_own, m, batch, true, e.s);
    }
}

// phase 2: ModUp to the owned targets, NTT, inner product with the owned key limbs, INTT of the owned P accumulators
void op_ks_shard_inner(OpEnv& e, const ShardPlan& sp, BatchPtr coeff_all, BatchPtr c_own_ntt, const u64* evk_own,
                       BatchPtr acc_own, BatchPtr accP_own, int batch, u64* const* peers) {
    Context& c = e.c;
    LSB_CUDA(cudaSetDevice(c.device));
    if (sp.n_own == 0) return;
    const size_t N = c.N;
    const size_t item = (size_t)sp.beta * sp.n_own * N;
    u64* scratch = e.ws.reserve(item * batch, e.s);
    BatchPtr ext = bp_strided(scratch, item);
    if (sp.n_up) launch_basis_extend(c.dt, sp.up_jobs, sp.n_up, coeff_all, ext, ext, batch, e.s, sp.up_meta.data());
    launch_ntt(c.dt, ext, ext, sp.ext_rows, batch, false, e.s);
    KsInnerArgs ia;
    ia.ext = ext;
    ia.cntt = c_own_ntt;
    ia.cntt_slot0 = 0;
    ia.evk = evk_own;
    ia.acc = acc_own;
    ia.l = sp.l;
    ia.k = sp.k;
    ia.beta = sp.beta;
    ia.alpha = sp.k;
    ia.key_level = sp.level;
    ia.nq = c.nq;
    ia.mapped = 1;
    ia.rows = sp.n_own;
    u64 qmax = 0;
    for (int u = 0; u < sp.nq_own; u++) {
        ia.tmap[u] = (uint8_t)sp.own_q[u];
        qmax = std::max(qmax, c.moduli[sp.own_q[u]]);
    }
    for (int u = 0; u < sp.np_own; u++) {
        ia.tmap[sp.nq_own + u] = (uint8_t)(sp.l + sp.own_p[u]);
        qmax = std::max(qmax, c.moduli[c.idxP(sp.own_p[u])]);
    }
    ia.fold = (int)std::min<u64>(1u << 20, ((1ull << 63) - 1) / qmax);
    ia.f64 = qmax < kElemF64Limit ? 1 : 0;
    launch_ks_inner(c.dt, ia, batch, e.s);
    if (sp.np_own > 0) {
        RowMap m;
        m.n = 0;
        for (int p = 0; p < 2; p++)
            for (int u = 0; u < sp.np_own; u++) {
                m.in_slot[m.n] = (uint8_t)(p * sp.n_own + sp.nq_own + u);
                m.out_slot[m.n] = (uint8_t)(peers ? (sp.rank * 2 + p) * sp.cp + u : p * sp.cp + u);
                m.mod[m.n] = (uint8_t)c.idxP(sp.own_p[u]);
                m.n++;
            }
        if (peers) {  // second all-gather fused into the INTT of the owned P accumulators (peers = every accP_all)
            LSB_CHECK(batch == 1, "lsb: the fused all-gather handles one ciphertext per call");
            launch_intt_to_peers(c.dt, acc_own, m, peers, sp.world, e.s);
        } else {
            launch_ntt(c.dt, acc_own, accP_own, m, batch, true, e.s);
        }
    }
}

// phase 3: ModDown from the gathered P limbs to the owned Q limbs, NTT + finish (+ addend_own when given)
void op_ks_shard_end(OpEnv& e, const ShardPlan& sp, BatchPtr accP_all, BatchPtr acc_own, BatchPtr addend_own, bool has_addend,
                     BatchPtr out_own, int batch) {
    Context& c = e.c;
    LSB_CUDA(cudaSetDevice(c.device));
    if (sp.nq_own == 0) return;
    const size_t N = c.N;
    const size_t item = (size_t)2 * sp.nq_own * N;
    u64* scratch = e.ws.reserve(item * batch, e.s);
    BatchPtr md = bp_strided(scratch, item);
    launch_basis_extend(c.dt, sp.down_jobs, sp.n_down, accP_all, md, md, batch, e.s, sp.down_meta.data());
    const int add = has_addend ? 1 : 0;
    NttFinish fin{sp.fin_dev[add], &sp.fin_host[add], acc_own, addend_own, out_own, has_addend};
    launch_ntt_finish(c.dt, md, fin, batch, e.s);
}

}  // namespace lsb
