// Per-(level, world, rank) tables of the limb-sharded key-switch (lsb_shard.cu).
#pragma once
#include <vector>
#include "lsb_kernels.h"

namespace lsb {

struct Context;

struct ShardPlan {
    int level, world, rank;
    int l, k, beta;
    int cq, cp;                  // padded owned-limb counts: ceil(l / world), ceil(k / world)
    std::vector<int> own_q, own_p;  // owned Q limb indices / P limb indices, ascending
    int nq_own, np_own, n_own;
    const BexJob* up_jobs = nullptr;
    int n_up = 0;
    std::vector<BexMeta> up_meta;
    RowMap ext_rows;             // forward NTT rows of ext_own [beta][n_own]
    const BexJob* down_jobs = nullptr;
    int n_down = 0;
    std::vector<BexMeta> down_meta;
    FinishMap fin_host[2];       // [0]: no addend, [1]: + addend_own
    const FinishMap* fin_dev[2] = {nullptr, nullptr};

    ShardPlan(Context& c, int level, int world, int rank);
};

}  // namespace lsb
