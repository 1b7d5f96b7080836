// TEST: the boundary types of include/lattisense_mega_ag.hpp / lattisense_gpu_runner.h against the REFERENCE's own headers,
// checked by the compiler (VERDICT r1 "weak" 4): both header sets are included in one translation unit -- the reference's
// inside namespace ref, with a stub for the cgo-generated liblattigo header -- and every member's size and offset is
// compared.  Built by tests/test_abi_layout_cpu.py with -DLATTISENSE_WITH_NLOHMANN_JSON -I<reference>/lib so that
// `attributes` is the same nlohmann::json on both sides.  Reference files: mega_ag_runners/mega_ag.h:43-177,
// mega_ag_runners/c_argument.h:24-46, abi/c_types.h:26-60, mega_ag_runners/wrapper.h:33-85.
#include <any>
#include <atomic>
#include <cmath>
#include <complex>
#include <cstddef>
#include <cstdint>
#include <functional>
#include <inttypes.h>
#include <map>
#include <memory>
#include <optional>
#include <set>
#include <stdbool.h>
#include <string>
#include <type_traits>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>
#include <gsl/span>
#include "nlohmann/json.hpp"

#include "lattisense_mega_ag.hpp"  // ours (global namespace), pulls lattisense_gpu_runner.h

namespace ref {
#include "mega_ag.h"      // reference (c_argument.h, fhe_lib_v2.h -> stub liblattigo)
#include "wrapper.h"
extern "C" {
#include "c_types.h"
}
}  // namespace ref

#define SAME_SIZE(T) static_assert(sizeof(::T) == sizeof(ref::T) && alignof(::T) == alignof(ref::T), "size/alignment of " #T)
#define SAME_OFF(T, m)                                                                                       \
    static_assert(offsetof(::T, m) == offsetof(ref::T, m) && sizeof(((::T*)nullptr)->m) == sizeof(((ref::T*)nullptr)->m), \
                  "offset/size of " #T "::" #m)

// ---- C wire structs (abi/c_types.h, c_argument.h) ----
SAME_SIZE(CComponent);
SAME_OFF(CComponent, n);
SAME_OFF(CComponent, data);
SAME_SIZE(CPolynomial);
SAME_OFF(CPolynomial, n_component);
SAME_OFF(CPolynomial, components);
SAME_SIZE(CPlaintext);
SAME_OFF(CPlaintext, level);
SAME_OFF(CPlaintext, poly);
SAME_SIZE(CCiphertext);
SAME_OFF(CCiphertext, level);
SAME_OFF(CCiphertext, degree);
SAME_OFF(CCiphertext, polys);
SAME_SIZE(CRelinKey);
SAME_SIZE(CKeySwitchKey);
SAME_SIZE(CGaloisKey);
SAME_OFF(CGaloisKey, n_key_switch_key);
SAME_OFF(CGaloisKey, galois_elements);
SAME_OFF(CGaloisKey, key_switch_keys);
SAME_SIZE(CArgument);
SAME_OFF(CArgument, id);
SAME_OFF(CArgument, type);
SAME_OFF(CArgument, data);
SAME_OFF(CArgument, level);
SAME_OFF(CArgument, size);
static_assert((int)::TYPE_PLAINTEXT == (int)ref::TYPE_PLAINTEXT && (int)::TYPE_CIPHERTEXT == (int)ref::TYPE_CIPHERTEXT &&
                  (int)::TYPE_RELIN_KEY == (int)ref::TYPE_RELIN_KEY && (int)::TYPE_GALOIS_KEY == (int)ref::TYPE_GALOIS_KEY &&
                  (int)::TYPE_SWITCH_KEY == (int)ref::TYPE_SWITCH_KEY && (int)::TYPE_CUSTOM == (int)ref::TYPE_CUSTOM,
              "DataType enumerators");
static_assert((int)::ALGO_BFV == (int)ref::ALGO_BFV && (int)::ALGO_CKKS == (int)ref::ALGO_CKKS, "Algo enumerators");

// ---- C++ types crossing the boundary (mega_ag.h) ----
SAME_SIZE(ExecutionContext);
SAME_OFF(ExecutionContext, context);
SAME_OFF(ExecutionContext, other_args);
static_assert(std::is_same<::NodeIndex, ref::NodeIndex>::value, "NodeIndex");
static_assert(sizeof(::ExecutorFunc) == sizeof(ref::ExecutorFunc), "ExecutorFunc");
#define SAME_ENUM(e) static_assert((int)::OperationType::e == (int)ref::OperationType::e, "OperationType::" #e)
SAME_ENUM(UNKNOWN);
SAME_ENUM(ADD);
SAME_ENUM(SUB);
SAME_ENUM(NEGATE);
SAME_ENUM(MULTIPLY);
SAME_ENUM(RELINEARIZE);
SAME_ENUM(RESCALE);
SAME_ENUM(DROP_LEVEL);
SAME_ENUM(ROTATE_COL);
SAME_ENUM(ROTATE_ROW);
SAME_ENUM(MAC_WO_PARTIAL_SUM);
SAME_ENUM(MAC_W_PARTIAL_SUM);
SAME_ENUM(BOOTSTRAP);
SAME_ENUM(FPGA_KERNEL);
SAME_ENUM(EXPORT_TO_ABI);
SAME_ENUM(IMPORT_FROM_ABI);
SAME_ENUM(LOAD_TO_BACKEND);
SAME_ENUM(STORE_FROM_BACKEND);
static_assert((int)::Processor::GPU == (int)ref::Processor::GPU && (int)::Processor::CPU == (int)ref::Processor::CPU, "Processor");

SAME_SIZE(DatumNode);
SAME_OFF(DatumNode, index);
SAME_OFF(DatumNode, id);
SAME_OFF(DatumNode, predecessors);
SAME_OFF(DatumNode, successors);
SAME_OFF(DatumNode, is_input);
SAME_OFF(DatumNode, is_output);
SAME_OFF(DatumNode, datum_type);
SAME_OFF(DatumNode, fhe_prop);
SAME_OFF(DatumNode, custom_prop);
SAME_SIZE(DatumNode::FheProperty);
SAME_OFF(DatumNode::FheProperty, level);
SAME_OFF(DatumNode::FheProperty, degree);
SAME_OFF(DatumNode::FheProperty, is_ntt);
SAME_OFF(DatumNode::FheProperty, is_mform);
SAME_OFF(DatumNode::FheProperty, p);
SAME_OFF(DatumNode::FheProperty, sp_level);
SAME_SIZE(DatumNode::FheProperty::ExtraProperty);
SAME_OFF(DatumNode::FheProperty::ExtraProperty, is_ringt);
SAME_OFF(DatumNode::FheProperty::ExtraProperty, is_compressed);
SAME_OFF(DatumNode::FheProperty::ExtraProperty, galois_element);
SAME_SIZE(DatumNode::CustomProperty);
SAME_OFF(DatumNode::CustomProperty, type);
SAME_OFF(DatumNode::CustomProperty, attributes);

SAME_SIZE(ComputeNode);
SAME_OFF(ComputeNode, index);
SAME_OFF(ComputeNode, id);
SAME_OFF(ComputeNode, input_nodes);
SAME_OFF(ComputeNode, output_nodes);
SAME_OFF(ComputeNode, executor);
SAME_OFF(ComputeNode, on_cpu);
SAME_OFF(ComputeNode, priority);
SAME_OFF(ComputeNode, sched_meta);
SAME_OFF(ComputeNode, fhe_prop);
SAME_OFF(ComputeNode, custom_prop);
SAME_SIZE(ComputeNode::ScheduleMeta);
SAME_OFF(ComputeNode::ScheduleMeta, top_level);
SAME_OFF(ComputeNode::ScheduleMeta, bottom_level);
SAME_SIZE(ComputeNode::FheProperty);
SAME_OFF(ComputeNode::FheProperty, op_type);
SAME_OFF(ComputeNode::FheProperty, p);
SAME_SIZE(ComputeNode::FheProperty::ExtraProperty);
SAME_OFF(ComputeNode::FheProperty::ExtraProperty, rotation_step);
SAME_OFF(ComputeNode::FheProperty::ExtraProperty, sum_cnt);
SAME_SIZE(ComputeNode::CustomProperty);
SAME_OFF(ComputeNode::CustomProperty, type);
SAME_OFF(ComputeNode::CustomProperty, attributes);

// ---- entry points: same prototypes (a mismatch fails to compile: the two declarations would differ only by namespace) ----
static_assert(std::is_same<decltype(&::run_fhe_gpu_task),
                           int (*)(::fhe_task_handle, ::CArgument*, uint64_t, ::CArgument*, uint64_t, ::progress_callback_t, void*, int)>::value,
              "run_fhe_gpu_task prototype");
static_assert(std::is_same<::progress_callback_t, ref::progress_callback_t>::value, "progress_callback_t");
static_assert(std::is_same<::fhe_task_handle, ref::fhe_task_handle>::value, "fhe_task_handle");

int main() { return 0; }
