// JSON graph model of the tachikoma runtime: JSONGraphNode / JSONGraphNodeEntry / JSONRuntimeBase::Load
// (src/runtime/contrib/json/json_node.h:51-324, json_runtime.h:181-260) re-implemented on json.h.
#pragma once
#include <array>
#include <cstdint>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "epilogue.cuh"
#include "json.h"

namespace tkb {

struct GNode {
  std::string op_type;  // "input" | "const" | "kernel"
  std::string name;     // variable / constant name, or operator (composite) name
  std::vector<std::array<uint32_t, 3>> inputs;  // (node id, output index, version)
  std::map<std::string, std::vector<std::string>> attrs;
  std::vector<std::vector<int64_t>> shapes;  // one per output
  std::vector<std::string> dtypes;
  uint32_t num_outputs = 1;

  bool has(const std::string& k) const { return attrs.count(k) != 0; }
  const std::vector<std::string>& list(const std::string& k) const {
    auto it = attrs.find(k);
    if (it == attrs.end()) throw std::runtime_error("Key: " + k + " is not found in node " + name);
    return it->second;
  }
  std::string str(const std::string& k, const std::string& dflt = "") const {
    auto it = attrs.find(k);
    return it == attrs.end() || it->second.empty() ? dflt : it->second[0];
  }
  int64_t i64(const std::string& k, int64_t dflt = 0) const {
    auto it = attrs.find(k);
    return it == attrs.end() || it->second.empty() || it->second[0].empty() ? dflt
                                                                             : std::strtoll(it->second[0].c_str(), nullptr, 10);
  }
  double f64(const std::string& k, double dflt = 0) const {
    auto it = attrs.find(k);
    return it == attrs.end() || it->second.empty() ? dflt : std::strtod(it->second[0].c_str(), nullptr);
  }
  std::vector<int64_t> ints(const std::string& k) const {
    std::vector<int64_t> out;
    auto it = attrs.find(k);
    if (it != attrs.end())
      for (auto& s : it->second) out.push_back(std::strtoll(s.c_str(), nullptr, 10));
    return out;
  }
};

struct Graph {
  std::vector<GNode> nodes;
  std::vector<uint32_t> arg_nodes, node_row_ptr;
  std::vector<std::array<uint32_t, 3>> heads;
  std::vector<uint32_t> input_eids;  // graph inputs, arg_nodes order
  std::vector<uint32_t> const_nids;  // constants, arg_nodes order

  uint32_t eid(uint32_t nid, uint32_t idx) const { return node_row_ptr[nid] + idx; }
  uint32_t eid(const std::array<uint32_t, 3>& e) const { return eid(e[0], e[1]); }
  uint32_t num_entries() const { return node_row_ptr.back(); }
};

inline std::array<uint32_t, 3> parse_entry(const JValue& v) {
  if (v.kind != JValue::Arr || v.arr.size() < 2) throw std::runtime_error("invalid json format: node entry");
  std::array<uint32_t, 3> e{(uint32_t)v.arr[0].as_int(), (uint32_t)v.arr[1].as_int(), 0};
  if (v.arr.size() > 2) e[2] = (uint32_t)v.arr[2].as_int();
  return e;
}

inline int dtype_from_string(const std::string& s) {
  if (s == "int8") return kI8;
  if (s == "uint8") return kU8;
  if (s == "int32") return kI32;
  if (s == "int16") return kI16;
  if (s == "int64") return kI64;
  if (s == "float32") return kF32;
  throw std::runtime_error("unsupported dtype: " + s);
}

inline Graph parse_graph(const std::string& json, const std::vector<std::string>& const_names) {
  JValue root = JParser(json).parse();
  Graph g;
  for (auto& kv : root.obj) {
    if (kv.first == "nodes") {
      for (auto& jn : kv.second.arr) {
        GNode n;
        for (auto& f : jn.obj) {
          if (f.first == "op") n.op_type = f.second.str;
          else if (f.first == "name") n.name = f.second.str;
          else if (f.first == "inputs") {
            for (auto& e : f.second.arr) n.inputs.push_back(parse_entry(e));
          } else if (f.first == "attr" || f.first == "attrs") {
            for (auto& a : f.second.obj) {
              if (a.first == "num_inputs") continue;
              if (a.first == "num_outputs") { n.num_outputs = (uint32_t)a.second.as_int(); continue; }
              if (a.second.kind != JValue::Arr || a.second.arr.size() != 1)
                throw std::runtime_error("invalid json format: attribute " + a.first);
              const JValue& inner = a.second.arr[0];
              if (a.first == "shape") {
                for (auto& s : inner.arr) {
                  std::vector<int64_t> shp;
                  for (auto& d : s.arr) shp.push_back(d.as_int());
                  n.shapes.push_back(shp);
                }
              } else if (a.first == "dtype") {
                for (auto& s : inner.arr) n.dtypes.push_back(s.str);
              } else {
                std::vector<std::string> vals;
                for (auto& s : inner.arr) vals.push_back(s.kind == JValue::Str ? s.str : std::to_string(s.num));
                n.attrs[a.first] = vals;
              }
            }
          } else {
            throw std::runtime_error("Unknown key: " + f.first);
          }
        }
        if (n.shapes.size() != n.dtypes.size()) throw std::runtime_error("shape/dtype count mismatch in " + n.name);
        g.nodes.push_back(std::move(n));
      }
    } else if (kv.first == "arg_nodes") {
      for (auto& v : kv.second.arr) g.arg_nodes.push_back((uint32_t)v.as_int());
    } else if (kv.first == "node_row_ptr") {
      for (auto& v : kv.second.arr) g.node_row_ptr.push_back((uint32_t)v.as_int());
    } else if (kv.first == "heads") {
      for (auto& v : kv.second.arr) g.heads.push_back(parse_entry(v));
    } else {
      throw std::runtime_error("Unknown key: " + kv.first);
    }
  }
  if (g.node_row_ptr.size() != g.nodes.size() + 1) throw std::runtime_error("node_row_ptr does not match nodes");
  // JSONRuntimeBase::LoadGraph (json_runtime.h:181-214): inputs and constants in arg_nodes order;
  // constants must appear in the same order as const_names.
  std::vector<std::string> consts;
  for (uint32_t nid : g.arg_nodes) {
    if (nid >= g.nodes.size()) throw std::runtime_error("arg_nodes out of range");
    GNode& n = g.nodes[nid];
    if (n.op_type == "input") {
      n.num_outputs = (uint32_t)n.shapes.size();
      for (uint32_t j = 0; j < n.shapes.size(); ++j) g.input_eids.push_back(g.eid(nid, j));
    } else if (n.op_type == "const") {
      bool found = false;
      for (auto& c : const_names) found = found || c == n.name;
      if (!found) throw std::runtime_error("Found non-existent constant: " + n.name);
      g.const_nids.push_back(nid);
      consts.push_back(n.name);
    } else {
      throw std::runtime_error("arg node " + n.name + " is neither input nor const");
    }
  }
  if (consts.size() != const_names.size())
    throw std::runtime_error("Found mismatch for the number of constants in the graph and required.");
  for (size_t i = 0; i < consts.size(); ++i)
    if (consts[i] != const_names[i])
      throw std::runtime_error("The position of constant in the graph must be the same as the required.");
  return g;
}

}  // namespace tkb
