// lattisense_mega_ag.hpp — the C++ types that cross the "C" boundary of libgpu_mega_ag_runner.so.
//
// The reference passes `ExecutorFunc*` through the void* parameters of
// bind_gpu_task_abi_bridge_executors / bind_gpu_task_custom_executors (mega_ag_runners/gpu/gpu_wrapper.cu:
// 492-510) and its executors read the graph nodes they are called with (e.g.
// cxx_sdk_v2/cxx_abi_bridge_executors.h:87-107: self.input_nodes[0]->datum_type / ->fhe_prop / ->is_input).
// For a caller compiled against the reference's own mega_ag_runners/mega_ag.h (:43-177) to interoperate,
// the node structs below keep the same members, in the same order, with the same types; the one
// member whose type belongs to a third-party library (CustomProperty::attributes, nlohmann::json in
// the reference) is last in its struct and is selected at build time:
//   -DLATTISENSE_WITH_NLOHMANN_JSON (with the reference's lib/ on the include path) -> nlohmann::json,
//   otherwise the built-in lsr::Json (what this repository's own tests and plug-in use).
// See INTEGRATION.md.
#pragma once
#include <any>
#include <cstdint>
#include <functional>
#include <optional>
#include <string>
#include <unordered_map>
#include <vector>

#include "lattisense_gpu_runner.h"

#ifdef LATTISENSE_WITH_NLOHMANN_JSON
#include "nlohmann/json.hpp"
namespace lsr { using AttrJson = nlohmann::json; }
#else
#include "lsr_json.h"
namespace lsr { using AttrJson = lsr::Json; }
#endif

using NodeIndex = uint64_t;

struct ComputeNode;

enum class Processor { CPU, FPGA, GPU };

// mega_ag.h:43-60.  On this backend: `context` is empty for GPU compute nodes and holds whatever the
// frontend-context factory returned for EXPORT/IMPORT/custom nodes (null when none is installed);
// other_args[0] of an IMPORT_FROM_ABI node is the caller's pre-allocated output handle (void*).
struct ExecutionContext {
    std::any context;
    std::vector<std::any> other_args;

    template <typename T> T* get_arithmetic_context() {
        auto* p = std::any_cast<T*>(&context);
        return p ? *p : nullptr;
    }
    template <typename T> T* get_other_arg(size_t index = 0) {
        if (index >= other_args.size() || !other_args[index].has_value()) return nullptr;
        return std::any_cast<T*>(other_args[index]);
    }
};

// mega_ag.h:63-66
using ExecutorFunc = std::function<void(ExecutionContext& ctx, const std::unordered_map<NodeIndex, std::any>& inputs,
                                        std::any& output, const ComputeNode& self)>;

// mega_ag.h:68-90 (same enumerator order: values are part of the layout contract)
enum class OperationType {
    UNKNOWN,
    ADD,
    SUB,
    NEGATE,
    MULTIPLY,
    RELINEARIZE,
    RESCALE,
    DROP_LEVEL,
    ROTATE_COL,
    ROTATE_ROW,
    MAC_WO_PARTIAL_SUM,
    MAC_W_PARTIAL_SUM,
    BOOTSTRAP,
    FPGA_KERNEL,
    EXPORT_TO_ABI,
    IMPORT_FROM_ABI,
    LOAD_TO_BACKEND,
    STORE_FROM_BACKEND,
};

// mega_ag.h:98-133
struct DatumNode {
    NodeIndex index;
    std::string id;
    std::vector<ComputeNode*> predecessors;
    std::vector<ComputeNode*> successors;
    bool is_input = false;
    bool is_output = false;
    DataType datum_type = TYPE_CUSTOM;

    struct FheProperty {
        int32_t level = 0;
        int32_t degree = 0;
        bool is_ntt = false;
        bool is_mform = false;
        struct ExtraProperty {
            bool is_ringt = false;
            bool is_compressed = false;
            uint32_t galois_element = 0;
        };
        std::optional<ExtraProperty> p;
        int32_t sp_level = 0;
    };
    std::optional<FheProperty> fhe_prop;

    struct CustomProperty {
        std::string type;
        lsr::AttrJson attributes;
    };
    std::optional<CustomProperty> custom_prop;
};

// mega_ag.h:138-177
struct ComputeNode {
    NodeIndex index;
    std::string id;
    std::vector<DatumNode*> input_nodes;
    std::vector<DatumNode*> output_nodes;
    ExecutorFunc executor;
    bool on_cpu = false;
    int priority = 0;
    struct ScheduleMeta {
        int top_level = 0;
        int bottom_level = 0;
    };
    ScheduleMeta sched_meta;

    struct FheProperty {
        OperationType op_type = OperationType::UNKNOWN;
        struct ExtraProperty {
            int32_t rotation_step = 0;
            int32_t sum_cnt = 0;
        };
        std::optional<ExtraProperty> p;
    };
    std::optional<FheProperty> fhe_prop;

    struct CustomProperty {
        std::string type;
        lsr::AttrJson attributes;
    };
    std::optional<CustomProperty> custom_prop;
};

// mega_ag_executors.h:53-54 (implementation mega_ag_executors_gpu.cu:463-512).  A caller that still links the reference's
// mega_ag.cpp (liblattisense.so, CMakeLists.txt:360-364) resolves this symbol while loading a graph.  This backend does
// not execute device nodes one by one -- run_fhe_gpu_task groups them into batched launches -- so the binder validates
// the (algorithm, operation) pair exactly as the reference does (same exceptions for unsupported pairs) and installs an
// executor that reports, if anybody calls it, that the node belongs to the batched dispatcher.
void bind_gpu_executor(ComputeNode& node, Algo algorithm);

// Optional hook (extension): EXPORT/IMPORT/custom executors of the in-tree SDK require an "empty frontend
// context" (BfvContext* / CkksContext* / CkksBtpContext*, gpu_wrapper.cu:175-177) in ExecutionContext::context.
// A caller that links liblattigo installs a factory producing that std::any from the task's `parameter`
// JSON text; plug-in executors ignore it.  Called once per run; the result is shared by all CPU nodes.
using FrontendContextFactory = std::function<std::any(Algo algo, const std::string& parameter_json)>;
extern "C" void lsb_gpu_task_set_frontend_context_factory(fhe_task_handle handle, void* factory /* FrontendContextFactory* */);
