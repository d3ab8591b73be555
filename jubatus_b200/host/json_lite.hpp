// Minimal JSON reader for the engine configs (config/nearest_neighbor/*.json,
// config/recommender/nearest_neighbor_recommender_*.json, config/anomaly/light_lof*.json).
// The reference parses these with jubatus::util::text::json + jsonconfig::config_cast_check
// (jubatus/server/server/nearest_neighbor_serv.cpp:82-93); only what that needs is here.
#ifndef JUBANN_HOST_JSON_LITE_HPP_
#define JUBANN_HOST_JSON_LITE_HPP_

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace jubatus_b200 {
namespace json {

struct value;
typedef std::shared_ptr<value> ptr;

struct value {
  enum kind_t { NUL, BOOL, NUM, STR, ARR, OBJ } kind = NUL;
  bool b = false;
  double num = 0;
  bool is_integer = false;
  std::string str;
  std::vector<ptr> arr;
  std::vector<std::pair<std::string, ptr> > obj;  // insertion order kept

  const ptr find(const std::string& k) const {
    for (size_t i = 0; i < obj.size(); ++i) if (obj[i].first == k) return obj[i].second;
    return ptr();
  }
};

class parse_error : public std::runtime_error {
 public:
  explicit parse_error(const std::string& m) : std::runtime_error("json syntax error: " + m) {}
};

class parser {
 public:
  explicit parser(const std::string& s) : s_(s), p_(0) {}
  ptr parse() {
    ptr v = value_();
    ws();
    if (p_ != s_.size()) throw parse_error("trailing characters");
    return v;
  }

 private:
  void ws() { while (p_ < s_.size() && (s_[p_] == ' ' || s_[p_] == '\n' || s_[p_] == '\t' || s_[p_] == '\r')) ++p_; }
  char peek() { ws(); if (p_ >= s_.size()) throw parse_error("unexpected end"); return s_[p_]; }
  void expect(char c) { if (peek() != c) throw parse_error(std::string("expected '") + c + "'"); ++p_; }
  ptr value_() {
    char c = peek();
    ptr v(new value);
    if (c == '{') {
      v->kind = value::OBJ; ++p_;
      if (peek() == '}') { ++p_; return v; }
      for (;;) {
        std::string k = string_();
        expect(':');
        v->obj.push_back(std::make_pair(k, value_()));
        if (peek() == ',') { ++p_; continue; }
        expect('}');
        return v;
      }
    } else if (c == '[') {
      v->kind = value::ARR; ++p_;
      if (peek() == ']') { ++p_; return v; }
      for (;;) {
        v->arr.push_back(value_());
        if (peek() == ',') { ++p_; continue; }
        expect(']');
        return v;
      }
    } else if (c == '"') {
      v->kind = value::STR; v->str = string_(); return v;
    } else if (s_.compare(p_, 4, "true") == 0) { p_ += 4; v->kind = value::BOOL; v->b = true; return v;
    } else if (s_.compare(p_, 5, "false") == 0) { p_ += 5; v->kind = value::BOOL; return v;
    } else if (s_.compare(p_, 4, "null") == 0) { p_ += 4; return v;
    } else {
      size_t b = p_;
      while (p_ < s_.size() && (isdigit(static_cast<unsigned char>(s_[p_])) || s_[p_] == '-' || s_[p_] == '+' ||
                                s_[p_] == '.' || s_[p_] == 'e' || s_[p_] == 'E')) ++p_;
      if (b == p_) throw parse_error("unexpected character");
      std::string t = s_.substr(b, p_ - b);
      char* end = 0;
      v->num = strtod(t.c_str(), &end);
      if (*end) throw parse_error("bad number " + t);
      v->kind = value::NUM;
      v->is_integer = t.find_first_of(".eE") == std::string::npos;
      return v;
    }
  }
  std::string string_() {
    expect('"');
    std::string out;
    while (p_ < s_.size() && s_[p_] != '"') {
      char c = s_[p_++];
      if (c == '\\') {
        if (p_ >= s_.size()) throw parse_error("bad escape");
        char e = s_[p_++];
        switch (e) {
          case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
          case 'b': out += '\b'; break; case 'f': out += '\f'; break;
          case 'u': {
            if (p_ + 4 > s_.size()) throw parse_error("bad \\u escape");
            unsigned cp = static_cast<unsigned>(strtoul(s_.substr(p_, 4).c_str(), 0, 16)); p_ += 4;
            if (cp < 0x80) out += static_cast<char>(cp);
            else if (cp < 0x800) { out += static_cast<char>(0xC0 | (cp >> 6)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
            else { out += static_cast<char>(0xE0 | (cp >> 12)); out += static_cast<char>(0x80 | ((cp >> 6) & 0x3F)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
            break;
          }
          default: out += e;
        }
      } else {
        out += c;
      }
    }
    expect('"');
    return out;
  }
  const std::string& s_;
  size_t p_;
};

inline ptr parse(const std::string& s) { return parser(s).parse(); }

// Canonical text of a value: no whitespace, object keys sorted (the reference's json object is a
// std::map, so re-serialising two configs that differ in spacing or key order gives equal text:
// compare_config, framework/save_load.cpp:102-109).
inline void dump_to(const value& v, std::string& out) {
  switch (v.kind) {
    case value::NUL: out += "null"; break;
    case value::BOOL: out += v.b ? "true" : "false"; break;
    case value::NUM: {
      char buf[40];
      if (v.is_integer) snprintf(buf, sizeof(buf), "%lld", static_cast<long long>(v.num));
      else snprintf(buf, sizeof(buf), "%.17g", v.num);
      out += buf;
      break;
    }
    case value::STR: {
      out.push_back('"');
      for (size_t i = 0; i < v.str.size(); ++i) {
        const unsigned char c = static_cast<unsigned char>(v.str[i]);
        if (c == '"' || c == '\\') { out.push_back('\\'); out.push_back(static_cast<char>(c)); }
        else if (c < 0x20) { char buf[8]; snprintf(buf, sizeof(buf), "\\u%04x", c); out += buf; }
        else out.push_back(static_cast<char>(c));
      }
      out.push_back('"');
      break;
    }
    case value::ARR:
      out.push_back('[');
      for (size_t i = 0; i < v.arr.size(); ++i) { if (i) out.push_back(','); dump_to(*v.arr[i], out); }
      out.push_back(']');
      break;
    case value::OBJ: {
      std::vector<size_t> order(v.obj.size());
      for (size_t i = 0; i < order.size(); ++i) order[i] = i;
      std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return v.obj[a].first < v.obj[b].first; });
      out.push_back('{');
      for (size_t i = 0; i < order.size(); ++i) {
        if (i) out.push_back(',');
        value k; k.kind = value::STR; k.str = v.obj[order[i]].first;
        dump_to(k, out);
        out.push_back(':');
        dump_to(*v.obj[order[i]].second, out);
      }
      out.push_back('}');
      break;
    }
  }
}
inline std::string dump(const ptr& v) { std::string out; if (v) dump_to(*v, out); return out; }

}  // namespace json
}  // namespace jubatus_b200
#endif
