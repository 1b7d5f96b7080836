// Task graph of one mega_ag.json: loader, ABI-bridge insertion for the GPU layout and scheduling levels.
// Semantics follow mega_ag_runners/mega_ag.cpp:125-657 (MegaAG::load = from_json + apply_processor_layout
// + compute_properties); node types are the boundary types of include/lattisense_mega_ag.hpp.
#pragma once
#include <string>
#include <unordered_map>
#include <vector>

#include "lattisense_mega_ag.hpp"
#include "lsr_json.h"

namespace lsr {

struct FheParameter {
    int n = 0;
    int max_level = 0;
    std::vector<uint64_t> q, p;
    uint64_t t = 0;  // BFV only
    bool has_btp = false;
};

struct TaskGraph {
    // node storage is address-stable (unordered_map nodes never move): nodes point at each other
    std::unordered_map<NodeIndex, DatumNode> data;
    std::unordered_map<NodeIndex, ComputeNode> computes;
    std::vector<NodeIndex> inputs, outputs, offline_inputs;
    Json parameter;
    FheParameter param;
    Algo algo = ALGO_BFV;
    std::string name;

    // mega_ag.cpp:125-130 with Processor::GPU
    static TaskGraph* load(const std::string& json_path);

    void bind_abi_bridge_executors(const ExecutorFunc& abi_export, const ExecutorFunc& abi_import);
    void bind_custom_executors(const std::unordered_map<std::string, ExecutorFunc>& custom);

    TaskGraph() = default;
    TaskGraph(const TaskGraph&) = delete;
    TaskGraph& operator=(const TaskGraph&) = delete;

private:
    void parse(const Json& root);
    void insert_gpu_bridge_nodes();
    void mark_cpu_nodes();
    void compute_levels();
};

// Serialise a Json value back to text (used to hand `parameter` to a frontend-context factory).
std::string json_dump(const Json& v);

inline bool is_key_type(DataType t) { return t == TYPE_RELIN_KEY || t == TYPE_GALOIS_KEY || t == TYPE_SWITCH_KEY; }

}  // namespace lsr
