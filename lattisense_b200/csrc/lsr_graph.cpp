// mega_ag.json -> TaskGraph (GPU layout).  See lsr_graph.h.
#include "lsr_graph.h"

#include <algorithm>
#include <deque>
#include <stdexcept>

namespace lsr {

// ---------------------------------------------------------------------------------------------
// JSON text output (compact)
// ---------------------------------------------------------------------------------------------
static void dump_into(const Json& v, std::string& out) {
    switch (v.kind) {
        case Json::Null: out += "null"; break;
        case Json::Bool: out += v.b ? "true" : "false"; break;
        case Json::UInt: out += std::to_string(v.u); break;
        case Json::Int: out += std::to_string(v.i); break;
        case Json::Real: {
            char buf[40];
            snprintf(buf, sizeof buf, "%.17g", v.d);
            out += buf;
            break;
        }
        case Json::String: {
            out.push_back('"');
            for (char c : v.s) {
                if (c == '"' || c == '\\') {
                    out.push_back('\\');
                    out.push_back(c);
                } else if (c == '\n') out += "\\n";
                else if (c == '\t') out += "\\t";
                else out.push_back(c);
            }
            out.push_back('"');
            break;
        }
        case Json::Array: {
            out.push_back('[');
            for (size_t i = 0; i < v.arr.size(); i++) {
                if (i) out.push_back(',');
                dump_into(v.arr[i], out);
            }
            out.push_back(']');
            break;
        }
        case Json::Object: {
            out.push_back('{');
            for (size_t i = 0; i < v.obj.size(); i++) {
                if (i) out.push_back(',');
                Json k;
                k.kind = Json::String;
                k.s = v.obj[i].first;
                dump_into(k, out);
                out.push_back(':');
                dump_into(v.obj[i].second, out);
            }
            out.push_back('}');
            break;
        }
    }
}
std::string json_dump(const Json& v) {
    std::string s;
    dump_into(v, s);
    return s;
}

// ---------------------------------------------------------------------------------------------
// type tables (mega_ag.cpp:27-47)
// ---------------------------------------------------------------------------------------------
static DataType datum_type_of(const std::string& t) {
    if (t == "ct" || t == "ct3") return TYPE_CIPHERTEXT;
    if (t == "pt" || t == "pt_mul" || t == "pt_ringt") return TYPE_PLAINTEXT;
    if (t == "rlk") return TYPE_RELIN_KEY;
    if (t == "glk") return TYPE_GALOIS_KEY;
    if (t == "swk") return TYPE_SWITCH_KEY;
    throw std::out_of_range("mega_ag: unknown datum type '" + t + "'");
}
static OperationType op_type_of(const std::string& t) {
    static const std::pair<const char*, OperationType> table[] = {
        {"add", OperationType::ADD},
        {"sub", OperationType::SUB},
        {"neg", OperationType::NEGATE},
        {"mult", OperationType::MULTIPLY},
        {"relin", OperationType::RELINEARIZE},
        {"rescale", OperationType::RESCALE},
        {"drop_level", OperationType::DROP_LEVEL},
        {"rotate_row", OperationType::ROTATE_ROW},
        {"rotate_col", OperationType::ROTATE_COL},
        {"cmp_sum", OperationType::MAC_WO_PARTIAL_SUM},
        {"cmpac_sum", OperationType::MAC_W_PARTIAL_SUM},
        {"bootstrap", OperationType::BOOTSTRAP},
        {"fpga_kernel", OperationType::FPGA_KERNEL},
    };
    for (auto& e : table)
        if (t == e.first) return e.second;
    throw std::out_of_range("mega_ag: unknown compute type '" + t + "'");
}

static bool flag(const Json& v, const char* key) {
    const Json* p = v.find(key);
    return p && p->kind == Json::Bool && p->b;
}

#ifdef LATTISENSE_WITH_NLOHMANN_JSON
static AttrJson to_attr(const Json& v) { return nlohmann::json::parse(json_dump(v)); }
#else
static AttrJson to_attr(const Json& v) { return v; }
#endif

// ---------------------------------------------------------------------------------------------
// from_json (mega_ag.cpp:132-305), GPU processor: key datum nodes are kept
// ---------------------------------------------------------------------------------------------
void TaskGraph::parse(const Json& root) {
    const std::string& algo_str = root.at("algorithm").as_string();
    if (algo_str == "BFV") algo = ALGO_BFV;
    else if (algo_str == "CKKS") algo = ALGO_CKKS;
    else throw std::runtime_error("Unknown algorithm: " + algo_str);
    if (const Json* nm = root.find("name"))
        if (nm->kind == Json::String) name = nm->s;

    for (auto& kv : root.at("data").obj) {
        const Json& v = kv.second;
        DatumNode node;
        node.index = std::stoull(kv.first);
        node.id = v.at("id").as_string();
        const std::string& jt = v.at("type").as_string();
        if (flag(v, "is_custom")) {
            DatumNode::CustomProperty cp;
            cp.type = jt;
            if (const Json* a = v.find("attributes")) cp.attributes = to_attr(*a);
            node.custom_prop = std::move(cp);
        } else {
            DatumNode::FheProperty fp;
            fp.level = v.at("level").as_int();
            fp.is_ntt = v.at("is_ntt").as_bool();
            fp.is_mform = v.at("is_mform").as_bool();
            fp.degree = v.at("degree").as_int();
            fp.sp_level = v.contains("sp_level") ? v.at("sp_level").as_int() : -1;
            node.datum_type = datum_type_of(jt);
            if (node.datum_type == TYPE_GALOIS_KEY) {
                DatumNode::FheProperty::ExtraProperty ep;
                ep.galois_element = (uint32_t)v.at("galois_element").as_u64();
                fp.p = ep;
            } else if (jt == "pt_ringt") {
                DatumNode::FheProperty::ExtraProperty ep;
                ep.is_ringt = true;
                fp.p = ep;
            }
            node.fhe_prop = fp;
        }
        NodeIndex idx = node.index;
        data.emplace(idx, std::move(node));
    }

    for (auto& kv : root.at("compute").obj) {
        const Json& v = kv.second;
        ComputeNode node;
        node.index = std::stoull(kv.first);
        node.id = v.at("id").as_string();
        const std::string& jt = v.at("type").as_string();
        if (flag(v, "is_custom")) {
            ComputeNode::CustomProperty cp;
            cp.type = jt;
            if (const Json* a = v.find("attributes")) cp.attributes = to_attr(*a);
            node.custom_prop = std::move(cp);
        } else {
            ComputeNode::FheProperty fp;
            fp.op_type = op_type_of(jt);
            if (fp.op_type == OperationType::ROTATE_COL) {
                ComputeNode::FheProperty::ExtraProperty ep;
                ep.rotation_step = v.at("step").as_int();
                fp.p = ep;
            } else if (fp.op_type == OperationType::MAC_WO_PARTIAL_SUM || fp.op_type == OperationType::MAC_W_PARTIAL_SUM) {
                ComputeNode::FheProperty::ExtraProperty ep;
                ep.sum_cnt = v.at("sum_cnt").as_int();
                fp.p = ep;
            }
            node.fhe_prop = fp;
        }
        for (uint64_t i : v.at("inputs").as_u64_vector()) node.input_nodes.push_back(&data.at(i));
        for (uint64_t i : v.at("outputs").as_u64_vector()) node.output_nodes.push_back(&data.at(i));
        NodeIndex idx = node.index;
        computes.emplace(idx, std::move(node));
    }
    for (auto& kv : computes) {
        for (DatumNode* d : kv.second.input_nodes) d->successors.push_back(&kv.second);
        for (DatumNode* d : kv.second.output_nodes) d->predecessors.push_back(&kv.second);
    }

    inputs = root.at("inputs").as_u64_vector();
    for (NodeIndex i : inputs) data.at(i).is_input = true;
    outputs = root.at("outputs").as_u64_vector();
    for (NodeIndex i : outputs) data.at(i).is_output = true;
    if (const Json* off = root.find("offline_inputs"))
        if (off->is_array()) offline_inputs = off->as_u64_vector();

    parameter = root.at("parameter");
    param.n = parameter.at("n").as_int();
    param.q = parameter.at("q").as_u64_vector();
    param.p = parameter.at("p").as_u64_vector();
    param.max_level = parameter.contains("max_level") ? parameter.at("max_level").as_int() : (int)param.q.size() - 1;
    // frontend/parameter.json:146-147 declares max_level 33 with 30 primes: clamp
    if (param.max_level > (int)param.q.size() - 1) param.max_level = (int)param.q.size() - 1;
    if (const Json* t = parameter.find("t"))
        if (t->is_number()) param.t = t->as_u64();
    param.has_btp = parameter.contains("btp_output_level");
}

// ---------------------------------------------------------------------------------------------
// ABI bridge insertion for the GPU layout (mega_ag.cpp:388-544).  A datum is "native" (a frontend
// handle) when it is a graph input or produced by a custom node; otherwise it lives on the device.
//   native datum with device consumers : d -> EXPORT_TO_ABI -> d_c_struct_h2d -> LOAD_TO_BACKEND -> d_gpu
//   device datum that is an output or feeds custom nodes :
//                         d_gpu -> STORE_FROM_BACKEND -> d_c_struct_d2h -> IMPORT_FROM_ABI -> d | d_handle
//   custom input datum                 : d -> EXPORT_TO_ABI -> d_concrete
//   native output produced by a custom node : d_concrete -> IMPORT_FROM_ABI -> d
// ---------------------------------------------------------------------------------------------
namespace {
struct BridgeBuilder {
    TaskGraph& g;
    NodeIndex next_datum = 0, next_compute = 0;

    explicit BridgeBuilder(TaskGraph& graph) : g(graph) {
        for (auto& kv : g.data) next_datum = std::max(next_datum, kv.first + 1);
        for (auto& kv : g.computes) next_compute = std::max(next_compute, kv.first + 1);
    }
    DatumNode& clone_datum(const DatumNode& src, const std::string& id) {
        DatumNode d = src;
        d.index = next_datum++;
        d.id = id;
        d.is_input = d.is_output = false;
        d.predecessors.clear();
        d.successors.clear();
        NodeIndex idx = d.index;
        return g.data.emplace(idx, std::move(d)).first->second;
    }
    ComputeNode& bridge(OperationType op, const std::string& id, DatumNode* in, DatumNode* out) {
        ComputeNode c;
        c.index = next_compute++;
        c.id = id;
        ComputeNode::FheProperty fp;
        fp.op_type = op;
        c.fhe_prop = fp;
        if (in) c.input_nodes.push_back(in);
        if (out) c.output_nodes.push_back(out);
        NodeIndex idx = c.index;
        return g.computes.emplace(idx, std::move(c)).first->second;
    }
    static void retarget_inputs(const std::vector<ComputeNode*>& consumers, DatumNode* from, DatumNode* to) {
        for (ComputeNode* c : consumers) {
            for (DatumNode*& in : c->input_nodes)
                if (in == from) in = to;
            to->successors.push_back(c);
        }
    }
};
}  // namespace

void TaskGraph::insert_gpu_bridge_nodes() {
    BridgeBuilder bb(*this);
    std::vector<NodeIndex> originals;
    originals.reserve(data.size());
    for (auto& kv : data) originals.push_back(kv.first);
    std::sort(originals.begin(), originals.end());  // deterministic numbering of the inserted nodes

    for (NodeIndex di : originals) {
        DatumNode& d = data.at(di);
        const std::string tag = std::to_string(di);
        bool native = true;
        if (!d.is_input) {
            if (d.predecessors.empty()) throw std::runtime_error("mega_ag: datum '" + d.id + "' has no producer and is not an input");
            native = d.predecessors[0]->custom_prop.has_value();
        }
        std::vector<ComputeNode*> dev_consumers, custom_consumers;
        for (ComputeNode* c : d.successors) (c->custom_prop.has_value() ? custom_consumers : dev_consumers).push_back(c);

        if (d.is_input && d.custom_prop.has_value()) {
            DatumNode& concrete = bb.clone_datum(d, d.id + "_concrete");
            bb.bridge(OperationType::EXPORT_TO_ABI, "export_to_abi_" + tag, &d, &concrete);
            std::vector<ComputeNode*> all = d.successors;
            d.successors.clear();
            BridgeBuilder::retarget_inputs(all, &d, &concrete);
            continue;
        }
        if (native && !dev_consumers.empty()) {
            DatumNode& cs = bb.clone_datum(d, d.id + "_c_struct_h2d");
            bb.bridge(OperationType::EXPORT_TO_ABI, "export_to_abi_" + tag, &d, &cs);
            DatumNode& dev = bb.clone_datum(cs, d.id + "_gpu");
            bb.bridge(OperationType::LOAD_TO_BACKEND, "load_to_gpu_" + tag, &cs, &dev);
            d.successors.clear();  // custom consumers of a native datum keep reading the handle (re-added below)
            BridgeBuilder::retarget_inputs(dev_consumers, &d, &dev);
            for (ComputeNode* c : custom_consumers) d.successors.push_back(c);
        }
        if (!native && (!custom_consumers.empty() || d.is_output)) {
            DatumNode& dev = bb.clone_datum(d, d.id + "_gpu");
            DatumNode& cs = bb.clone_datum(d, d.id + "_c_struct_d2h");
            bb.bridge(OperationType::STORE_FROM_BACKEND, "store_from_gpu_" + tag, &dev, &cs);
            ComputeNode& imp = bb.bridge(OperationType::IMPORT_FROM_ABI, "import_from_abi_" + tag, &cs, nullptr);
            if (d.is_output) {
                imp.output_nodes.push_back(&d);
            } else {
                DatumNode& handle = bb.clone_datum(cs, d.id + "_handle");
                imp.output_nodes.push_back(&handle);
                BridgeBuilder::retarget_inputs(custom_consumers, &d, &handle);
            }
            // device producers now write the device datum; device consumers read it
            for (ComputeNode* prod : d.predecessors) {
                if (!prod->fhe_prop.has_value()) continue;
                for (DatumNode*& out : prod->output_nodes)
                    if (out == &d) out = &dev;
                dev.predecessors.push_back(prod);
            }
            d.predecessors.clear();
            d.successors.clear();
            BridgeBuilder::retarget_inputs(dev_consumers, &d, &dev);
        }
        if (native && d.is_output && !d.is_input) {
            DatumNode& concrete = bb.clone_datum(d, d.id + "_concrete");
            bb.bridge(OperationType::IMPORT_FROM_ABI, "import_from_abi_" + tag, &concrete, &d);
            for (ComputeNode* prod : d.predecessors) {
                for (DatumNode*& out : prod->output_nodes)
                    if (out == &d) out = &concrete;
                concrete.predecessors.push_back(prod);
            }
            d.predecessors.clear();
        }
    }
    // link the bridge nodes themselves
    for (auto& kv : computes) {
        ComputeNode& c = kv.second;
        if (!c.fhe_prop.has_value()) continue;
        const OperationType op = c.fhe_prop->op_type;
        if (op != OperationType::EXPORT_TO_ABI && op != OperationType::IMPORT_FROM_ABI && op != OperationType::LOAD_TO_BACKEND &&
            op != OperationType::STORE_FROM_BACKEND)
            continue;
        for (DatumNode* in : c.input_nodes) in->successors.push_back(&c);
        for (DatumNode* out : c.output_nodes) out->predecessors.push_back(&c);
    }
}

// mega_ag.cpp:307-334 (GPU branch): custom nodes and EXPORT/IMPORT run on host threads
void TaskGraph::mark_cpu_nodes() {
    for (auto& kv : computes) {
        ComputeNode& c = kv.second;
        if (c.custom_prop.has_value()) c.on_cpu = true;
        else if (c.fhe_prop.has_value())
            c.on_cpu = c.fhe_prop->op_type == OperationType::EXPORT_TO_ABI || c.fhe_prop->op_type == OperationType::IMPORT_FROM_ABI;
        else c.on_cpu = false;
    }
}

// mega_ag.cpp:336-350, 587-657: longest path from a source (top) / to a sink (bottom); priority = bottom level
void TaskGraph::compute_levels() {
    std::unordered_map<NodeIndex, int> pending;
    std::deque<ComputeNode*> ready;
    for (auto& kv : computes) {
        ComputeNode& c = kv.second;
        c.sched_meta.top_level = c.sched_meta.bottom_level = 0;
        int deg = 0;
        for (DatumNode* in : c.input_nodes) deg += (int)in->predecessors.size();
        pending[c.index] = deg;
        if (deg == 0) ready.push_back(&c);
    }
    size_t visited = 0;
    while (!ready.empty()) {
        ComputeNode* u = ready.front();
        ready.pop_front();
        visited++;
        for (DatumNode* out : u->output_nodes)
            for (ComputeNode* v : out->successors) {
                v->sched_meta.top_level = std::max(v->sched_meta.top_level, u->sched_meta.top_level + 1);
                if (--pending[v->index] == 0) ready.push_back(v);
            }
    }
    if (visited != computes.size()) throw std::runtime_error("mega_ag: the compute graph has a cycle");
    for (auto& kv : computes) {
        ComputeNode& c = kv.second;
        int deg = 0;
        for (DatumNode* out : c.output_nodes) deg += (int)out->successors.size();
        pending[c.index] = deg;
        if (deg == 0) ready.push_back(&c);
    }
    while (!ready.empty()) {
        ComputeNode* v = ready.front();
        ready.pop_front();
        for (DatumNode* in : v->input_nodes)
            for (ComputeNode* u : in->predecessors) {
                u->sched_meta.bottom_level = std::max(u->sched_meta.bottom_level, v->sched_meta.bottom_level + 1);
                if (--pending[u->index] == 0) ready.push_back(u);
            }
    }
    for (auto& kv : computes) kv.second.priority = kv.second.sched_meta.bottom_level;
}

TaskGraph* TaskGraph::load(const std::string& json_path) {
    std::unique_ptr<TaskGraph> g(new TaskGraph());
    Json root;
    try {
        root = Json::parse_file(json_path);
    } catch (const std::runtime_error& e) {
        if (std::string(e.what()).rfind("Cannot open", 0) == 0) throw std::runtime_error("Cannot open MegaAG file " + json_path);
        throw;
    }
    g->parse(root);
    g->insert_gpu_bridge_nodes();
    g->mark_cpu_nodes();
    g->compute_levels();
    return g.release();
}

void TaskGraph::bind_abi_bridge_executors(const ExecutorFunc& abi_export, const ExecutorFunc& abi_import) {
    for (auto& kv : computes) {
        ComputeNode& c = kv.second;
        if (!c.fhe_prop.has_value()) continue;
        if (c.fhe_prop->op_type == OperationType::EXPORT_TO_ABI) c.executor = abi_export;
        else if (c.fhe_prop->op_type == OperationType::IMPORT_FROM_ABI) c.executor = abi_import;
    }
}
void TaskGraph::bind_custom_executors(const std::unordered_map<std::string, ExecutorFunc>& custom) {
    for (auto& kv : computes) {
        ComputeNode& c = kv.second;
        if (!c.custom_prop.has_value()) continue;
        auto it = custom.find(c.custom_prop->type);
        if (it != custom.end()) c.executor = it->second;
    }
}

}  // namespace lsr
