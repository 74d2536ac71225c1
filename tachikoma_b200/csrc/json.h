// Minimal JSON reader for the tachikoma graph schema (replaces dmlc::JSONReader use in
// src/runtime/contrib/json/json_node.h:99-324 and json_runtime.h:227-244).  Header only.
#pragma once
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace tkb {

struct JValue {
  enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
  bool b = false;
  double num = 0;
  std::string str;
  std::vector<JValue> arr;
  std::vector<std::pair<std::string, JValue>> obj;  // insertion order kept

  const JValue* get(const std::string& key) const {
    for (auto& kv : obj)
      if (kv.first == key) return &kv.second;
    return nullptr;
  }
  const JValue& at(const std::string& key) const {
    const JValue* v = get(key);
    if (!v) throw std::runtime_error("JSON: missing key '" + key + "'");
    return *v;
  }
  int64_t as_int() const {
    if (kind == Num) return (int64_t)num;
    if (kind == Str) return std::strtoll(str.c_str(), nullptr, 10);
    throw std::runtime_error("JSON: expected a number");
  }
};

class JParser {
 public:
  explicit JParser(const std::string& s) : s_(s) {}
  JValue parse() {
    JValue v = value();
    ws();
    if (p_ != s_.size()) fail("trailing characters");
    return v;
  }

 private:
  const std::string& s_;
  size_t p_ = 0;
  [[noreturn]] void fail(const char* m) {
    throw std::runtime_error(std::string("JSON parse error at ") + std::to_string(p_) + ": " + m);
  }
  void ws() {
    while (p_ < s_.size() && (s_[p_] == ' ' || s_[p_] == '\n' || s_[p_] == '\t' || s_[p_] == '\r')) ++p_;
  }
  JValue value() {
    ws();
    if (p_ >= s_.size()) fail("unexpected end");
    char c = s_[p_];
    JValue v;
    if (c == '{') {
      v.kind = JValue::Obj;
      ++p_;
      ws();
      if (s_[p_] == '}') { ++p_; return v; }
      for (;;) {
        ws();
        if (s_[p_] != '"') fail("expected key");
        std::string k = string();
        ws();
        if (s_[p_] != ':') fail("expected ':'");
        ++p_;
        v.obj.emplace_back(k, value());
        ws();
        if (s_[p_] == ',') { ++p_; continue; }
        if (s_[p_] == '}') { ++p_; break; }
        fail("expected ',' or '}'");
      }
    } else if (c == '[') {
      v.kind = JValue::Arr;
      ++p_;
      ws();
      if (s_[p_] == ']') { ++p_; return v; }
      for (;;) {
        v.arr.push_back(value());
        ws();
        if (s_[p_] == ',') { ++p_; continue; }
        if (s_[p_] == ']') { ++p_; break; }
        fail("expected ',' or ']'");
      }
    } else if (c == '"') {
      v.kind = JValue::Str;
      v.str = string();
    } else if (c == 't' && s_.compare(p_, 4, "true") == 0) {
      v.kind = JValue::Bool; v.b = true; p_ += 4;
    } else if (c == 'f' && s_.compare(p_, 5, "false") == 0) {
      v.kind = JValue::Bool; v.b = false; p_ += 5;
    } else if (c == 'n' && s_.compare(p_, 4, "null") == 0) {
      p_ += 4;
    } else {
      char* end = nullptr;
      v.kind = JValue::Num;
      v.num = std::strtod(s_.c_str() + p_, &end);
      if (end == s_.c_str() + p_) fail("bad value");
      p_ = end - s_.c_str();
    }
    return v;
  }
  std::string string() {
    std::string out;
    ++p_;  // opening quote
    while (p_ < s_.size() && s_[p_] != '"') {
      char c = s_[p_++];
      if (c == '\\') {
        if (p_ >= s_.size()) fail("bad escape");
        char e = s_[p_++];
        switch (e) {
          case 'n': out += '\n'; break;
          case 't': out += '\t'; break;
          case 'r': out += '\r'; break;
          case 'b': out += '\b'; break;
          case 'f': out += '\f'; break;
          case 'u': {
            unsigned cp = (unsigned)std::strtoul(s_.substr(p_, 4).c_str(), nullptr, 16);
            p_ += 4;
            if (cp < 0x80) out += (char)cp;
            else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
            else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
            break;
          }
          default: out += e;
        }
      } else {
        out += c;
      }
    }
    if (p_ >= s_.size()) fail("unterminated string");
    ++p_;
    return out;
  }
};

}  // namespace tkb
