// Minimal JSON reader for cameras.json (the reference uses nlohmann/json via FetchContent,
// CMakeLists.txt:48-57; only objects / arrays / numbers / strings / literals are needed).
#pragma once
#include <cctype>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>
namespace rt::scene::json_lite {

class Value {
 public:
  enum Kind { kNull, kBool, kNumber, kString, kArray, kObject };
  Kind kind = kNull;
  double num = 0;
  bool flag = false;
  std::string str;
  std::vector<Value> items;
  std::vector<std::pair<std::string, Value>> fields;  // insertion order kept

  double number() const {
    if (kind != kNumber) throw std::runtime_error("json: number expected");
    return num;
  }
  const Value& at(size_t i) const {
    if (kind != kArray || i >= items.size()) throw std::runtime_error("json: array index out of range");
    return items[i];
  }
  const Value* find(const std::string& key) const {
    for (const auto& f : fields) if (f.first == key) return &f.second;
    return nullptr;
  }
  const Value& at(const std::string& key) const {
    const Value* v = kind == kObject ? find(key) : nullptr;
    if (!v) throw std::runtime_error("json: missing key '" + key + "'");
    return *v;
  }
  double number_or(const std::string& key, double fallback) const {
    const Value* v = kind == kObject ? find(key) : nullptr;
    return (v && v->kind == kNumber) ? v->num : fallback;
  }
  const std::vector<std::pair<std::string, Value>>& members() const {
    if (kind != kObject) throw std::runtime_error("json: object expected");
    return fields;
  }
};

class Parser {
 public:
  explicit Parser(const std::string& s) : s_(s) {}
  Value Run() {
    Value v = ParseValue();
    Skip();
    if (i_ != s_.size()) Fail("trailing characters");
    return v;
  }
 private:
  const std::string& s_;
  size_t i_ = 0;
  [[noreturn]] void Fail(const std::string& why) const {
    throw std::runtime_error("json: " + why + " at offset " + std::to_string(i_));
  }
  void Skip() { while (i_ < s_.size() && std::isspace(static_cast<unsigned char>(s_[i_]))) ++i_; }
  bool Eat(char c) { Skip(); if (i_ < s_.size() && s_[i_] == c) { ++i_; return true; } return false; }
  std::string ParseString() {
    if (!Eat('"')) Fail("'\"' expected");
    std::string out;
    while (i_ < s_.size() && s_[i_] != '"') {
      char c = s_[i_++];
      if (c == '\\' && i_ < s_.size()) {
        char e = s_[i_++];
        switch (e) { case 'n': c = '\n'; break; case 't': c = '\t'; break; case 'r': c = '\r'; break;
                     case 'b': c = '\b'; break; case 'f': c = '\f'; break; default: c = e; }
      }
      out.push_back(c);
    }
    if (i_ >= s_.size()) Fail("unterminated string");
    ++i_;
    return out;
  }
  Value ParseValue() {
    Skip();
    if (i_ >= s_.size()) Fail("unexpected end");
    Value v;
    const char c = s_[i_];
    if (c == '{') {
      ++i_; v.kind = Value::kObject;
      if (Eat('}')) return v;
      do {
        std::string key = ParseString();
        if (!Eat(':')) Fail("':' expected");
        v.fields.emplace_back(std::move(key), ParseValue());
      } while (Eat(','));
      if (!Eat('}')) Fail("'}' expected");
    } else if (c == '[') {
      ++i_; v.kind = Value::kArray;
      if (Eat(']')) return v;
      do { v.items.push_back(ParseValue()); } while (Eat(','));
      if (!Eat(']')) Fail("']' expected");
    } else if (c == '"') {
      v.kind = Value::kString; v.str = ParseString();
    } else if (s_.compare(i_, 4, "true") == 0) { v.kind = Value::kBool; v.flag = true; i_ += 4;
    } else if (s_.compare(i_, 5, "false") == 0) { v.kind = Value::kBool; i_ += 5;
    } else if (s_.compare(i_, 4, "null") == 0) { i_ += 4;
    } else {
      const char* begin = s_.c_str() + i_;
      char* end = nullptr;
      v.num = std::strtod(begin, &end);
      if (end == begin) Fail("value expected");
      v.kind = Value::kNumber;
      i_ += static_cast<size_t>(end - begin);
    }
    return v;
  }
};

inline Value Parse(const std::string& text) { return Parser(text).Run(); }
inline std::string ReadFile(const std::string& path) {
  std::ifstream f(path);
  if (!f) throw std::runtime_error("cannot open '" + path + "'");
  std::stringstream ss;
  ss << f.rdbuf();
  return ss.str();
}
}  // namespace rt::scene::json_lite
