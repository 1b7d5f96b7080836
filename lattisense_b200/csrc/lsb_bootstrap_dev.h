// Device-side CKKS bootstrapping (the `bootstrap` node, mega_ag_executors_gpu.cu:410-426): see lsb_bootstrap.h for the
// algorithm and lsb_bootstrap_dev.cu for the device operations underneath.
#pragma once
#include <memory>
#include <vector>
#include "lsb_bootstrap.h"
#include "lsb_ops.h"

namespace lsb {

// device pointers of the keys a bootstrap node receives: [rlk, glk..., swk_dts, swk_std] (custom_task.py:1961-1996)
struct BootstrapKeys {
    const u64* rlk = nullptr;
    std::vector<u64> gal;           // Galois element of glk[i] (column rotations and 2N-1 for the row key)
    std::vector<const u64*> glk;
    const u64* swk_dts = nullptr;   // dense -> sparse secret, level swk_dts_level (0)
    const u64* swk_std = nullptr;   // sparse -> dense secret, level key_level
    int key_level = 0, swk_dts_level = 0;
};
struct BootstrapStats {
    int rotations = 0, key_switches = 0, mults = 0, rescales = 0, plain_mults = 0;
};

class DeviceBootstrap {
public:
    DeviceBootstrap(Context& c, const btp::Params& p);
    ~DeviceBootstrap();
    DeviceBootstrap(const DeviceBootstrap&) = delete;
    DeviceBootstrap& operator=(const DeviceBootstrap&) = delete;
    // in: batch of level-0 ciphertexts [2][1][N]; out: [2][out_level+1][N]; asynchronous on `s`
    void run(Workspace& ws, cudaStream_t s, const BootstrapKeys& keys, BatchPtr in, BatchPtr out, int batch, BootstrapStats* st = nullptr);
    std::vector<int> rotations() const;
    int out_level() const;

private:
    struct Impl;
    Context& ctx_;
    std::unique_ptr<Impl> impl_;
};

}  // namespace lsb
