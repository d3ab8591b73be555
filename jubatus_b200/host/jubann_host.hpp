// Host side of the drop-in: the reference's C++ interfaces for the nearest-neighbor path,
// re-stated over the C-ABI of libjubann.so (include/jubann.h).
//
// What mirrors what (jubatus_core 0.3.2 is un-vendored; cited by the reference call site):
//   jubatus_b200::nearest_neighbor_base / gpu_nearest_neighbor / create_nearest_neighbor
//       <- core::nearest_neighbor::nearest_neighbor_base, {lsh,euclid_lsh,minhash},
//          create_nearest_neighbor   (jubatus/server/server/nearest_neighbor_serv.cpp:106-108)
//   jubatus_b200::device_table       <- core::storage::column_table (keys, versions, index on
//          the host; columns on the device)          (nearest_neighbor_serv.cpp:99-100)
//   jubatus_b200::mixable_versioned_table <- core::framework::mixable_versioned_table
//          (driven by framework/mixer/push_mixer.cpp:410-472, linear_mixer.cpp:547-561,613-628)
//   jubatus_b200::driver::nearest_neighbor <- core::driver::nearest_neighbor
//          (nearest_neighbor_serv.cpp:125,134,144,153,162,171,177)
//   jubatus_b200::nearest_neighbor_recommender <- core::recommender::nearest_neighbor_recommender
//          (recommender_serv.cpp:111-116,136-144,170-200)
//   jubatus_b200::light_lof          <- core::anomaly::light_lof
//          (anomaly_serv.cpp:134-138,169,216,224,238)
//   jubatus_b200::datum_to_fv_converter <- core::fv_converter (num / str rules of the shipped
//          configs only, config/nearest_neighbor/lsh.json:2-15)
//   jubatus_b200::nearest_neighbor_serv <- jubatus::server::nearest_neighbor_serv (RPC surface of
//          nearest_neighbor.idl:6-27, config schema of nearest_neighbor_serv.cpp:45-54)
// Errors are exceptions, as in the reference (JUBATUS_EXCEPTION); the RPC dispatcher turns them
// into error strings (common/mprpc/rpc_server.cpp:41-53).
#ifndef JUBANN_HOST_HPP_
#define JUBANN_HOST_HPP_

#include <ctype.h>
#include <dlfcn.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <fstream>
#include <functional>
#include <list>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <thread>
#include <unordered_set>
#include <utility>
#include <vector>

#include "jubann.h"
#include "json_lite.hpp"
#include "model_file.hpp"

namespace jubatus_b200 {

// ---------------------------------------------------------------------------------------
// exceptions (names follow jubatus::core::common::exception)
// ---------------------------------------------------------------------------------------
class jubatus_exception : public std::runtime_error {
 public:
  explicit jubatus_exception(const std::string& m) : std::runtime_error(m) {}
};
class config_exception : public jubatus_exception {
 public:
  explicit config_exception(const std::string& m) : jubatus_exception("config_exception: " + m) {}
};
class invalid_parameter : public jubatus_exception {
 public:
  explicit invalid_parameter(const std::string& m) : jubatus_exception("invalid_parameter: " + m) {}
};
class unsupported_method : public jubatus_exception {
 public:
  explicit unsupported_method(const std::string& m) : jubatus_exception("unsupported_method: " + m) {}
};
class converter_exception : public jubatus_exception {
 public:
  explicit converter_exception(const std::string& m) : jubatus_exception("converter_exception: " + m) {}
};

typedef std::vector<std::pair<std::string, float> > sfv_t;  // core::common::sfv_t

// jubatus::client::common::datum wire struct (jubatus/client/common/datum.hpp:29-47)
struct datum {
  std::vector<std::pair<std::string, std::string> > string_values;
  std::vector<std::pair<std::string, double> > num_values;
  std::vector<std::pair<std::string, std::string> > binary_values;
  datum& add_string(const std::string& k, const std::string& v) { string_values.push_back(std::make_pair(k, v)); return *this; }
  datum& add_number(const std::string& k, double v) { num_values.push_back(std::make_pair(k, v)); return *this; }
};

// common::hash_util::calc_string_hash: 64-bit FNV-1, `char` sign-extends
inline uint64_t calc_string_hash(const std::string& s) {
  uint64_t h = 14695981039346656037ULL;
  for (size_t i = 0; i < s.size(); ++i) {
    h *= 1099511628211ULL;
    h ^= static_cast<uint64_t>(static_cast<int64_t>(static_cast<signed char>(s[i])));
  }
  return h;
}

// ---------------------------------------------------------------------------------------
// libjubann.so, loaded like an fv_converter plugin
// (jubatus/server/fv_converter/dynamic_loader.cpp:44-94: absolute/relative path as given,
//  else $JUBATUS_PLUGIN_PATH/<name>, else the configured plugin dir)
// ---------------------------------------------------------------------------------------
struct jubann_api {
  void* dl;
  const char* (*version)(void);
  const char* (*last_error)(void);
  int (*create)(const char*, uint32_t, int, uint32_t, uint32_t, jubann_t**);
  int (*destroy)(jubann_t*);
  int (*set_option)(jubann_t*, const char*, int64_t);
  int (*clear)(jubann_t*);
  int (*size)(jubann_t*, uint64_t*, uint64_t*);
  int (*hash_csr)(jubann_t*, uint64_t, const uint64_t*, const uint32_t*, const uint64_t*, const float*, uint64_t*, float*);
  int (*set_rows_csr)(jubann_t*, uint64_t, const uint64_t*, const uint64_t*, const uint32_t*, const uint64_t*, const float*);
  int (*set_rows_sig)(jubann_t*, uint64_t, const uint64_t*, const uint64_t*, const float*);
  int (*get_rows_sig)(jubann_t*, uint64_t, const uint64_t*, uint64_t*, float*);
  int (*move_row)(jubann_t*, uint64_t, uint64_t);
  int (*truncate)(jubann_t*, uint64_t);
  int (*query_sig)(jubann_t*, uint64_t, const uint64_t*, const float*, uint32_t, uint32_t, uint64_t*, float*, uint32_t*);
  int (*query_csr)(jubann_t*, uint64_t, const uint64_t*, const uint32_t*, const uint64_t*, const float*, uint32_t, uint32_t, uint64_t*, float*, uint32_t*);
  int (*query_rows)(jubann_t*, uint64_t, const uint64_t*, uint32_t, uint32_t, uint64_t*, float*, uint32_t*);
  // more than one GPU
  int (*create_multi)(const char*, uint32_t, const int*, int, jubann_t**);
  int (*comm_unique_id)(void*);
  int (*comm_init)(jubann_t*, const void*, uint32_t, uint32_t);
  int (*comm_info)(jubann_t*, uint32_t*, uint32_t*);
  int (*mix_allgather)(jubann_t*, uint64_t, const uint64_t*, const void*, uint64_t, uint64_t*, uint64_t*, const void**);
  int (*mix_apply)(jubann_t*, uint64_t, const uint32_t*, const uint32_t*, const uint64_t*);
  // optional (0 when the library predates it): set_row + neighbor_row(id) in one call
  int (*set_row_query_csr)(jubann_t*, uint64_t, uint64_t, const uint32_t*, const uint64_t*, const float*, uint32_t, uint32_t,
                           uint64_t*, float*, uint32_t*);

  static std::shared_ptr<jubann_api> load(const std::string& path_or_name = "libjubann.so") {
    void* dl = 0;
    std::string tried;
    if (path_or_name.find('/') != std::string::npos) {
      dl = ::dlopen(path_or_name.c_str(), RTLD_LAZY);
      tried = path_or_name;
    } else {
      const char* dir = ::getenv("JUBATUS_PLUGIN_PATH");
      if (dir) { tried = std::string(dir) + "/" + path_or_name; dl = ::dlopen(tried.c_str(), RTLD_LAZY); }
      if (!dl) { tried = path_or_name; dl = ::dlopen(path_or_name.c_str(), RTLD_LAZY); }
    }
    if (!dl) {
      const char* e = ::dlerror();
      throw jubatus_exception("cannot load dynamic library: " + tried + ": " + (e ? e : "?"));
    }
    std::shared_ptr<jubann_api> a(new jubann_api);
    a->dl = dl;
#define JUBANN_SYM(field, name)                                                     \
  *reinterpret_cast<void**>(&a->field) = ::dlsym(dl, name);                          \
  if (!a->field) throw jubatus_exception(std::string("symbol not found: ") + name)
    JUBANN_SYM(version, "version");
    JUBANN_SYM(last_error, "jubann_last_error");
    JUBANN_SYM(create, "jubann_create");
    JUBANN_SYM(destroy, "jubann_destroy");
    JUBANN_SYM(set_option, "jubann_set_option");
    JUBANN_SYM(clear, "jubann_clear");
    JUBANN_SYM(size, "jubann_size");
    JUBANN_SYM(hash_csr, "jubann_hash_csr");
    JUBANN_SYM(set_rows_csr, "jubann_set_rows_csr");
    JUBANN_SYM(set_rows_sig, "jubann_set_rows_sig");
    JUBANN_SYM(get_rows_sig, "jubann_get_rows_sig");
    JUBANN_SYM(move_row, "jubann_move_row");
    JUBANN_SYM(truncate, "jubann_truncate");
    JUBANN_SYM(query_sig, "jubann_query_sig");
    JUBANN_SYM(query_csr, "jubann_query_csr");
    JUBANN_SYM(query_rows, "jubann_query_rows");
    JUBANN_SYM(create_multi, "jubann_create_multi");
    JUBANN_SYM(comm_unique_id, "jubann_comm_unique_id");
    JUBANN_SYM(comm_init, "jubann_comm_init");
    JUBANN_SYM(comm_info, "jubann_comm_info");
    JUBANN_SYM(mix_allgather, "jubann_mix_allgather");
    JUBANN_SYM(mix_apply, "jubann_mix_apply");
#undef JUBANN_SYM
    *reinterpret_cast<void**>(&a->set_row_query_csr) = ::dlsym(dl, "jubann_set_row_query_csr");
    return a;
  }
  void check(int rc) const {
    if (rc == 0) return;
    const std::string msg = last_error();
    if (rc == JUBANN_E_INVALID) {
      if (msg.compare(0, 18, "unsupported_method") == 0) throw unsupported_method(msg.substr(20));
      throw invalid_parameter(msg);
    }
    throw jubatus_exception(msg);
  }
};

// ---------------------------------------------------------------------------------------
// weight_manager -- global feature weights ([core] fv_converter/weight_manager, keyword_weights;
// restated from recall).  A feature whose key ends in "/idf" is scaled by
// log((documents + 1) / (documents_with_the_feature + 1)); "/bin" and everything else by 1.
// Counts move only in convert_and_update_weight (set_row, update_row, add, train); queries read
// them.  Features that end up with value 0 are dropped.  Counts are kept as (master, diff) so
// that replicas can exchange diffs (get_diff / mix / put_diff, like the reference's mixable).
// ---------------------------------------------------------------------------------------
class weight_manager {
 public:
  struct counts {
    uint64_t documents;
    std::unordered_map<std::string, uint64_t> frequency;
    counts() : documents(0) {}
    void add(const counts& o) {
      documents += o.documents;
      for (std::unordered_map<std::string, uint64_t>::const_iterator it = o.frequency.begin(); it != o.frequency.end(); ++it)
        frequency[it->first] += it->second;
    }
  };
  void update_weight(const sfv_t& fv) {
    ++diff_.documents;
    for (size_t i = 0; i < fv.size(); ++i)
      if (is_idf(fv[i].first)) diff_.frequency[fv[i].first] += 1;
  }
  double get_global_weight(const std::string& key) const {
    if (!is_idf(key)) return 1.0;
    const double n = static_cast<double>(master_.documents + diff_.documents);
    const double df = static_cast<double>(lookup(master_, key) + lookup(diff_, key));
    return log((n + 1.0) / (df + 1.0));
  }
  void get_weight(sfv_t& fv) const {
    size_t w = 0;
    for (size_t i = 0; i < fv.size(); ++i) {
      const float v = static_cast<float>(fv[i].second * get_global_weight(fv[i].first));
      if (v != 0.0f) { if (w != i) fv[w].first.swap(fv[i].first); fv[w].second = v; ++w; }
    }
    fv.resize(w);
  }
  void clear() { master_ = counts(); diff_ = counts(); }
  void add_documents(uint64_t n) { diff_.documents += n; }   // n stored rows without any idf feature
  // MIX: every replica's diff is summed and the sum lands in every master
  void get_diff(counts& out) const { out = diff_; }
  static void mix(const counts& lhs, counts& acc) { acc.add(lhs); }
  void put_diff(const counts& mixed) { master_.add(mixed); diff_ = counts(); }
  uint64_t documents() const { return master_.documents + diff_.documents; }
  // a diff on the wire (MIX)
  static std::string pack_counts(const counts& c) {
    std::string b;
    put64(b, c.documents);
    put64(b, c.frequency.size());
    for (std::unordered_map<std::string, uint64_t>::const_iterator it = c.frequency.begin(); it != c.frequency.end(); ++it) {
      put64(b, it->first.size()); b += it->first; put64(b, it->second);
    }
    return b;
  }
  static counts unpack_counts(const std::string& b) {
    size_t p = 0;
    counts c;
    c.documents = get64(b, p);
    const uint64_t n = get64(b, p);
    for (uint64_t i = 0; i < n; ++i) {
      const uint64_t len = get64(b, p);
      if (b.size() < p + len) throw jubatus_exception("unpack: truncated weights");
      std::string key(b, p, len);
      p += len;
      c.frequency[key] = get64(b, p);
    }
    return c;
  }
  // save / load: master and diff folded together (a loaded model starts with an empty diff)
  std::string pack() const {
    counts all = master_;
    all.add(diff_);
    std::string b;
    put64(b, all.documents);
    put64(b, all.frequency.size());
    std::vector<std::string> keys;
    for (std::unordered_map<std::string, uint64_t>::const_iterator it = all.frequency.begin(); it != all.frequency.end(); ++it) keys.push_back(it->first);
    std::sort(keys.begin(), keys.end());
    for (size_t i = 0; i < keys.size(); ++i) { put64(b, keys[i].size()); b += keys[i]; put64(b, all.frequency[keys[i]]); }
    return b;
  }
  void unpack(const std::string& b) {
    size_t p = 0;
    counts c;
    c.documents = get64(b, p);
    const uint64_t n = get64(b, p);
    for (uint64_t i = 0; i < n; ++i) {
      const uint64_t len = get64(b, p);
      if (b.size() < p + len) throw jubatus_exception("unpack: truncated weights");
      std::string key(b, p, len);
      p += len;
      c.frequency[key] = get64(b, p);
    }
    master_ = c;
    diff_ = counts();
  }
 private:
  static void put64(std::string& b, uint64_t v) { char t[8]; memcpy(t, &v, 8); b.append(t, 8); }
  static uint64_t get64(const std::string& b, size_t& p) {
    if (b.size() < p + 8) throw jubatus_exception("unpack: truncated weights");
    uint64_t v; memcpy(&v, b.data() + p, 8); p += 8; return v;
  }
  static bool is_idf(const std::string& key) { return key.size() >= 4 && key.compare(key.size() - 4, 4, "/idf") == 0; }
  static uint64_t lookup(const counts& c, const std::string& key) {
    std::unordered_map<std::string, uint64_t>::const_iterator it = c.frequency.find(key);
    return it == c.frequency.end() ? 0 : it->second;
  }
  counts master_, diff_;
};

// ---------------------------------------------------------------------------------------
// datum -> sfv ([core] fv_converter/datum_to_fv_converter; the rule kinds the shipped configs of
// this path use: config/nearest_neighbor/*.json, config/recommender/nearest_neighbor_recommender*.json,
// config/anomaly/light_lof*.json, config/classifier/nn.json).
//   string rules: type "str" (the whole value), "space" (split at white space) or a name from
//     "string_types" with {"method": "ngram", "char_num": "N"} (character N-grams, UTF-8 aware);
//     sample_weight bin | tf | log_tf; global_weight bin | idf.  Key:
//     <key>$<token>@<type>#<sample_weight>/<global_weight>; tokens of one value in sorted order
//     (the reference counts them in a std::map).
//   num rules: num | log | str.      combination rules: mul | add.      hash_max_size.
// ---------------------------------------------------------------------------------------
class datum_to_fv_converter {
 public:
  struct rule { std::string key, type, sample_weight, global_weight; int ngram; bool space; };
  struct combination_rule { std::string key_left, key_right, type; };
  datum_to_fv_converter() : hash_max_size_(0) {}

  static std::shared_ptr<datum_to_fv_converter> from_config(const json::ptr& conv) {
    std::shared_ptr<datum_to_fv_converter> c(new datum_to_fv_converter);
    if (!conv || conv->kind != json::value::OBJ) throw config_exception("converter must be an object");
    static const char* const known[] = {"string_filter_types", "string_filter_rules", "num_filter_types",
                                        "num_filter_rules", "string_types", "string_rules", "num_types", "num_rules",
                                        "binary_types", "binary_rules", "combination_types", "combination_rules",
                                        "hash_max_size"};
    for (size_t i = 0; i < conv->obj.size(); ++i) {
      bool ok = false;
      for (size_t j = 0; j < sizeof(known) / sizeof(known[0]); ++j) ok = ok || conv->obj[i].first == known[j];
      if (!ok) throw config_exception("unknown converter key: " + conv->obj[i].first);
    }
    // string_types: name -> ngram width (the only splitter method built here)
    std::map<std::string, int> ngram_types;
    json::ptr st = conv->find("string_types");
    if (st) for (size_t i = 0; i < st->obj.size(); ++i) {
      const json::ptr& t = st->obj[i].second;
      json::ptr m = t ? t->find("method") : json::ptr();
      if (!m || m->kind != json::value::STR) throw converter_exception("string type needs a \"method\": " + st->obj[i].first);
      if (m->str != "ngram") throw converter_exception("string type method not supported by this build: " + m->str);
      json::ptr n = t->find("char_num");
      int width = 0;
      if (n && n->kind == json::value::STR) width = atoi(n->str.c_str());
      else if (n && n->kind == json::value::NUM) width = static_cast<int>(n->num);
      if (width <= 0) throw converter_exception("ngram needs a positive \"char_num\": " + st->obj[i].first);
      ngram_types[st->obj[i].first] = width;
    }
    json::ptr nr = conv->find("num_rules");
    if (nr) for (size_t i = 0; i < nr->arr.size(); ++i) {
      rule r; r.ngram = 0; r.space = false; r.key = str(nr->arr[i], "key"); r.type = str(nr->arr[i], "type");
      if (r.type != "num" && r.type != "log" && r.type != "str") throw converter_exception("unknown num feature type: " + r.type);
      c->num_rules_.push_back(r);
    }
    json::ptr sr = conv->find("string_rules");
    if (sr) for (size_t i = 0; i < sr->arr.size(); ++i) {
      rule r; r.ngram = 0; r.space = false; r.key = str(sr->arr[i], "key"); r.type = str(sr->arr[i], "type");
      r.sample_weight = str(sr->arr[i], "sample_weight"); r.global_weight = str(sr->arr[i], "global_weight");
      if (r.type == "space") r.space = true;
      else if (ngram_types.count(r.type)) r.ngram = ngram_types[r.type];
      else if (r.type != "str") throw converter_exception("unknown string feature type: " + r.type);
      if (r.sample_weight != "bin" && r.sample_weight != "tf" && r.sample_weight != "log_tf") throw converter_exception("unknown sample_weight: " + r.sample_weight);
      if (r.global_weight != "bin" && r.global_weight != "idf") throw converter_exception("global_weight not supported by this build: " + r.global_weight);
      c->string_rules_.push_back(r);
    }
    // combination features (config/anomaly/light_lof_combinational_feature.json:15-18): "mul" and
    // "add" of every pair of converted features whose keys match (key_left, key_right) in either
    // order; [core] fv_converter/combination_feature*, restated from recall: pairs i < j in fv
    // order, key "<left>&<right>/<type>"
    json::ptr cr = conv->find("combination_rules");
    if (cr) for (size_t i = 0; i < cr->arr.size(); ++i) {
      combination_rule r;
      r.key_left = str(cr->arr[i], "key_left"); r.key_right = str(cr->arr[i], "key_right"); r.type = str(cr->arr[i], "type");
      if (r.type != "mul" && r.type != "add") throw converter_exception("unknown combination feature type: " + r.type);
      c->combination_rules_.push_back(r);
    }
    // sections this build has no implementation for must be empty: ignoring a rule would change
    // every signature silently
    static const char* const unsupported[] = {"string_filter_types", "string_filter_rules", "num_filter_types",
                                              "num_filter_rules", "num_types", "binary_types", "binary_rules",
                                              "combination_types"};
    for (size_t j = 0; j < sizeof(unsupported) / sizeof(unsupported[0]); ++j) {
      json::ptr v = conv->find(unsupported[j]);
      if (v && (!v->arr.empty() || !v->obj.empty()))
        throw converter_exception(std::string("converter section not supported by this build: ") + unsupported[j]);
    }
    json::ptr hm = conv->find("hash_max_size");
    if (hm && hm->kind == json::value::NUM) c->hash_max_size_ = static_cast<uint64_t>(hm->num);
    return c;
  }

  // datum_to_fv_converter::convert: queries -- global weights are read, not moved
  void convert(const datum& d, sfv_t& out) const {
    convert_unweighted(d, out);
    weights_.get_weight(out);
    finish(out);
  }
  // datum_to_fv_converter::convert_and_update_weight: everything that stores a row
  void convert_and_update_weight(const datum& d, sfv_t& out) {
    convert_unweighted(d, out);
    weights_.update_weight(out);
    weights_.get_weight(out);
    finish(out);
  }
  // A batch of stored rows, same result as convert_and_update_weight row by row.  Without an idf
  // rule a row's weights do not depend on the rows before it (only the document count moves), so
  // the conversion itself is spread over `threads` host threads; with one, rows go in order.
  bool has_idf() const {
    for (size_t r = 0; r < string_rules_.size(); ++r) if (string_rules_[r].global_weight == "idf") return true;
    return false;
  }
  template <typename MakeDatum>
  void convert_and_update_weight_batch(size_t n, MakeDatum make, std::vector<sfv_t>& out, unsigned threads) {
    out.resize(n);
    if (has_idf() || threads < 2 || n < 1024) {
      for (size_t i = 0; i < n; ++i) convert_and_update_weight(make(i), out[i]);
      return;
    }
    std::vector<std::thread> th;
    for (unsigned t = 0; t < threads; ++t)
      th.push_back(std::thread([&, t]() {
        for (size_t i = n * t / threads; i < n * (t + 1) / threads; ++i) {
          convert_unweighted(make(i), out[i]);
          weights_.get_weight(out[i]);
          finish(out[i]);
        }
      }));
    for (size_t t = 0; t < th.size(); ++t) th[t].join();
    weights_.add_documents(n);
  }
  weight_manager& weights() { return weights_; }
  const weight_manager& weights() const { return weights_; }
  void clear_weights() { weights_.clear(); }

 private:
  // strings first, then numbers, each in datum order and rule order
  void convert_unweighted(const datum& d, sfv_t& out) const {
    out.clear();
    for (size_t r = 0; r < string_rules_.size(); ++r) {
      const rule& sr = string_rules_[r];
      const std::string suffix = "@" + sr.type + "#" + sr.sample_weight + "/" + sr.global_weight;
      if (sr.ngram == 0 && !sr.space) {   // "str": the value is the token
        std::map<std::string, float> tf;
        std::vector<std::string> order;
        for (size_t i = 0; i < d.string_values.size(); ++i) {
          if (!match(sr.key, d.string_values[i].first)) continue;
          std::string f = d.string_values[i].first + "$" + d.string_values[i].second + suffix;
          if (tf.find(f) == tf.end()) order.push_back(f);
          tf[f] += 1.0f;
        }
        for (size_t i = 0; i < order.size(); ++i) out.push_back(std::make_pair(order[i], sample_weight(sr, tf[order[i]])));
        continue;
      }
      for (size_t i = 0; i < d.string_values.size(); ++i) {
        if (!match(sr.key, d.string_values[i].first)) continue;
        std::map<std::string, float> counter;   // sorted, like the reference's word counter
        if (sr.space) split_space(d.string_values[i].second, counter); else split_ngram(d.string_values[i].second, sr.ngram, counter);
        for (std::map<std::string, float>::const_iterator it = counter.begin(); it != counter.end(); ++it)
          out.push_back(std::make_pair(d.string_values[i].first + "$" + it->first + suffix, sample_weight(sr, it->second)));
      }
    }
    for (size_t r = 0; r < num_rules_.size(); ++r) {
      for (size_t i = 0; i < d.num_values.size(); ++i) {
        if (!match(num_rules_[r].key, d.num_values[i].first)) continue;
        const std::string& t = num_rules_[r].type;
        if (t == "num") out.push_back(std::make_pair(d.num_values[i].first + "@num", static_cast<float>(d.num_values[i].second)));
        else if (t == "log") out.push_back(std::make_pair(d.num_values[i].first + "@log", static_cast<float>(log(std::max(1.0, d.num_values[i].second)))));
        else {
          char buf[64]; snprintf(buf, sizeof(buf), "%g", d.num_values[i].second);
          out.push_back(std::make_pair(d.num_values[i].first + "$" + buf + "@str", 1.0f));
        }
      }
    }
    if (!combination_rules_.empty()) {
      const size_t n = out.size();
      for (size_t r = 0; r < combination_rules_.size(); ++r) {
        const combination_rule& cr = combination_rules_[r];
        for (size_t i = 0; i < n; ++i)
          for (size_t j = i + 1; j < n; ++j) {
            const bool fwd = match(cr.key_left, out[i].first) && match(cr.key_right, out[j].first);
            const bool rev = match(cr.key_left, out[j].first) && match(cr.key_right, out[i].first);
            if (!fwd && !rev) continue;
            const float v = cr.type == "mul" ? out[i].second * out[j].second : out[i].second + out[j].second;
            out.push_back(std::make_pair(out[i].first + "&" + out[j].first + "/" + cr.type, v));
          }
      }
    }
  }
  void finish(sfv_t& out) const {
    if (hash_max_size_) {
      for (size_t i = 0; i < out.size(); ++i) out[i].first = std::to_string(calc_string_hash(out[i].first) % hash_max_size_);
    }
  }
  static float sample_weight(const rule& r, float tf) {
    if (r.sample_weight == "bin") return 1.0f;
    if (r.sample_weight == "tf") return tf;
    return static_cast<float>(log(1.0 + tf));   // log_tf
  }
  static size_t utf8_len(unsigned char c) { return c < 0x80 ? 1 : (c >> 5) == 0x6 ? 2 : (c >> 4) == 0xE ? 3 : (c >> 3) == 0x1E ? 4 : 1; }
  static void split_ngram(const std::string& s, int n, std::map<std::string, float>& counter) {
    std::vector<size_t> starts;   // byte offsets of the characters, plus the end
    for (size_t p = 0; p < s.size(); p += utf8_len(static_cast<unsigned char>(s[p]))) starts.push_back(p);
    starts.push_back(s.size());
    const size_t chars = starts.size() - 1;
    for (size_t i = 0; i + n <= chars; ++i) counter[s.substr(starts[i], std::min(starts[i + n], s.size()) - starts[i])] += 1.0f;
  }
  static void split_space(const std::string& s, std::map<std::string, float>& counter) {
    size_t p = 0;
    while (p < s.size()) {
      while (p < s.size() && isspace(static_cast<unsigned char>(s[p]))) ++p;
      size_t e = p;
      while (e < s.size() && !isspace(static_cast<unsigned char>(s[e]))) ++e;
      if (e > p) counter[s.substr(p, e - p)] += 1.0f;
      p = e;
    }
  }
  static std::string str(const json::ptr& o, const char* k) {
    json::ptr v = o ? o->find(k) : json::ptr();
    if (!v || v->kind != json::value::STR) throw config_exception(std::string("rule needs a string \"") + k + "\"");
    return v->str;
  }
  static bool match(const std::string& pat, const std::string& key) {
    if (pat == "*" || pat.empty()) return true;
    if (pat[0] == '*') { std::string suf = pat.substr(1); return key.size() >= suf.size() && key.compare(key.size() - suf.size(), suf.size(), suf) == 0; }
    if (pat[pat.size() - 1] == '*') return key.compare(0, pat.size() - 1, pat, 0, pat.size() - 1) == 0;
    return pat == key;
  }
  std::vector<rule> num_rules_, string_rules_;
  std::vector<combination_rule> combination_rules_;
  uint64_t hash_max_size_;
  weight_manager weights_;
};

// ---------------------------------------------------------------------------------------
// device_table: column_table with the columns on the GPU
// ---------------------------------------------------------------------------------------
typedef std::pair<std::string, uint64_t> version_t;       // (owner, clock)
typedef std::map<std::string, uint64_t> version_clock;     // owner -> newest clock seen

struct packed_row {                                        // what column_table::get_row serialises
  std::string key;
  version_t version;
  std::vector<uint64_t> sig;
  float norm;
};

class device_table {
 public:
  device_table(const std::shared_ptr<jubann_api>& api, const std::string& method, uint32_t hash_num, int device)
      : api_(api), h_(0), hash_num_(hash_num), w64_((hash_num + 63) / 64), clock_(0), multi_(false) {
    api_->check(api_->create(method.c_str(), hash_num, device, 0, 1, &h_));
  }
  // one server process, several GPUs: the rows are sharded round-robin over `devices` inside the
  // library (jubann_create_multi); everything above this class is unchanged
  device_table(const std::shared_ptr<jubann_api>& api, const std::string& method, uint32_t hash_num,
               const std::vector<int>& devices)
      : api_(api), h_(0), hash_num_(hash_num), w64_((hash_num + 63) / 64), clock_(0), multi_(devices.size() > 1) {
    if (devices.size() == 1) api_->check(api_->create(method.c_str(), hash_num, devices[0], 0, 1, &h_));
    else api_->check(api_->create_multi(method.c_str(), hash_num, devices.data(), static_cast<int>(devices.size()), &h_));
  }
  ~device_table() { if (h_) api_->destroy(h_); }
  device_table(const device_table&) = delete;
  device_table& operator=(const device_table&) = delete;

  jubann_t* handle() const { return h_; }
  bool multi_device() const { return multi_; }
  const jubann_api& api() const { return *api_; }
  uint64_t size() const { return keys_.size(); }
  uint32_t words() const { return w64_; }
  void clear() {
    api_->check(api_->clear(h_));
    keys_.clear(); versions_.clear(); index_.clear(); clock_ = 0; dirty_.clear();
  }
  // rows stamped locally since the last MIX round (what get_diff has to look at: a row whose
  // version is newer than the agreed clocks got that version here, through assign or unpack)
  std::vector<uint64_t> dirty_rows() const {
    std::vector<uint64_t> d(dirty_.begin(), dirty_.end());
    std::sort(d.begin(), d.end());
    return d;
  }
  void clear_dirty() { dirty_.clear(); }
  std::pair<bool, uint64_t> exact_match(const std::string& key) const {
    std::unordered_map<std::string, uint64_t>::const_iterator it = index_.find(key);
    if (it == index_.end()) return std::make_pair(false, 0);
    return std::make_pair(true, it->second);
  }
  const std::string& get_key(uint64_t idx) const { return keys_.at(idx); }
  const version_t& get_version(uint64_t idx) const { return versions_.at(idx); }
  std::vector<std::string> get_all_row_ids() const { return keys_; }

  // column_table::add for a batch: assigns row indices (append or in-place), stamps
  // (owner, clock++) per row, returns the indices.  `undo` (optional) records what was there
  // before, so that a device call that fails afterwards leaves no row the table does not hold.
  struct assign_journal {
    uint64_t rows_before, clock_before;
    std::vector<std::pair<uint64_t, version_t> > overwritten;
  };
  std::vector<uint64_t> assign(const std::vector<std::string>& keys, const std::string& owner, assign_journal* undo = 0) {
    std::vector<uint64_t> idx(keys.size());
    if (undo) { undo->rows_before = keys_.size(); undo->clock_before = clock_; undo->overwritten.clear(); }
    for (size_t i = 0; i < keys.size(); ++i) {
      std::unordered_map<std::string, uint64_t>::iterator it = index_.find(keys[i]);
      if (it == index_.end()) {
        idx[i] = keys_.size();
        keys_.push_back(keys[i]);
        versions_.push_back(version_t(owner, clock_));
        index_[keys[i]] = idx[i];
      } else {
        idx[i] = it->second;
        if (undo && idx[i] < undo->rows_before) undo->overwritten.push_back(std::make_pair(idx[i], versions_[idx[i]]));
        versions_[idx[i]] = version_t(owner, clock_);
      }
      dirty_.insert(idx[i]);
      ++clock_;
    }
    return idx;
  }
  void rollback(const assign_journal& j) {
    while (keys_.size() > j.rows_before) { index_.erase(keys_.back()); keys_.pop_back(); versions_.pop_back(); dirty_.erase(keys_.size()); }
    for (size_t i = j.overwritten.size(); i-- > 0;) versions_[j.overwritten[i].first] = j.overwritten[i].second;
    clock_ = j.clock_before;
  }
  // column_table::delete_row: the last row moves into the hole
  bool delete_row(const std::string& key) {
    std::unordered_map<std::string, uint64_t>::iterator it = index_.find(key);
    if (it == index_.end()) return false;
    const uint64_t idx = it->second, last = keys_.size() - 1;
    if (idx != last) {
      api_->check(api_->move_row(h_, last, idx));
      keys_[idx] = keys_[last];
      versions_[idx] = versions_[last];
      index_[keys_[idx]] = idx;
      if (dirty_.count(last)) dirty_.insert(idx); else dirty_.erase(idx);
    } else {
      dirty_.erase(idx);
    }
    dirty_.erase(last);
    api_->check(api_->truncate(h_, last));
    index_.erase(key);
    keys_.pop_back(); versions_.pop_back();
    return true;
  }
  packed_row get_row(uint64_t idx) const {
    packed_row r;
    r.key = keys_.at(idx); r.version = versions_[idx]; r.sig.resize(w64_); r.norm = 0;
    api_->check(api_->get_rows_sig(h_, 1, &idx, r.sig.data(), &r.norm));
    return r;
  }
  void get_rows(const std::vector<uint64_t>& idx, std::vector<packed_row>& out) const {
    out.resize(idx.size());
    if (idx.empty()) return;
    std::vector<uint64_t> sigs(idx.size() * w64_);
    std::vector<float> norms(idx.size());
    api_->check(api_->get_rows_sig(h_, idx.size(), idx.data(), sigs.data(), norms.data()));
    for (size_t i = 0; i < idx.size(); ++i) {
      out[i].key = keys_.at(idx[i]); out[i].version = versions_[idx[i]];
      out[i].sig.assign(sigs.begin() + i * w64_, sigs.begin() + (i + 1) * w64_);
      out[i].norm = norms[i];
    }
  }
  // column_table::set_row(packed) for a batch: upsert by key; an existing row is overwritten
  // only if its clock <= the incoming clock; the local clock moves past every incoming clock.
  void set_rows(const std::vector<packed_row>& rows, std::vector<version_t>* applied_versions) {
    std::vector<uint64_t> idx, sigs;
    std::vector<float> norms;
    for (size_t i = 0; i < rows.size(); ++i) {
      const packed_row& r = rows[i];
      std::unordered_map<std::string, uint64_t>::iterator it = index_.find(r.key);
      bool store = true;
      uint64_t at;
      if (it == index_.end()) {
        at = keys_.size();
        keys_.push_back(r.key); versions_.push_back(r.version); index_[r.key] = at;
      } else {
        at = it->second;
        if (versions_[at].second <= r.version.second) versions_[at] = r.version; else store = false;
      }
      if (clock_ <= r.version.second) clock_ = r.version.second + 1;
      if (applied_versions) applied_versions->push_back(r.version);
      if (store) {
        idx.push_back(at);
        sigs.insert(sigs.end(), r.sig.begin(), r.sig.end());
        norms.push_back(r.norm);
      }
    }
    // rows are appended in increasing index order; an index may repeat, and the library stores the
    // last entry for it (what applying the batch in order gives)
    if (!idx.empty()) api_->check(api_->set_rows_sig(h_, idx.size(), idx.data(), sigs.data(), norms.data()));
  }
  uint64_t clock() const { return clock_; }
  // the host half of column_table::set_row(packed) for one incoming (key, version): unknown key ->
  // new row at the end; known key -> admitted only if the stored clock <= the incoming one.  The
  // local clock moves past every clock seen.  *at = the row the caller has to store into.
  bool admit(const std::string& key, const version_t& v, uint64_t* at) {
    std::unordered_map<std::string, uint64_t>::iterator it = index_.find(key);
    bool store = true;
    if (it == index_.end()) {
      *at = keys_.size();
      keys_.push_back(key); versions_.push_back(v); index_[key] = *at;
    } else {
      *at = it->second;
      if (versions_[*at].second <= v.second) versions_[*at] = v; else store = false;
    }
    if (clock_ <= v.second) clock_ = v.second + 1;
    return store;
  }

  // column_table::dump_json as the server's get_status uses it (nearest_neighbor_serv.cpp:72-73):
  // a short description, not the rows.
  std::string dump_json() const {
    return "{\"rows\":" + std::to_string(keys_.size()) + ",\"hash_num\":" + std::to_string(hash_num_) +
           ",\"clock\":" + std::to_string(clock_) + ",\"columns\":[\"bit_vector(" + std::to_string(hash_num_) +
           ")\"],\"storage\":\"device\"}";
  }

  // column_table::pack / unpack (driver->pack / unpack, framework/save_load.cpp:129,280).  The
  // reference writes msgpack; this is a self-describing little-endian blob with the same content
  // (keys, versions, clock, columns) -- not byte-compatible with a .jubatus model file.
  std::string pack() const {
    std::string out;
    put_u64(out, 0x314E4E4142554Aull);   // "JUBANN1"
    put_u64(out, hash_num_); put_u64(out, clock_); put_u64(out, keys_.size());
    std::vector<uint64_t> idx(keys_.size());
    for (size_t i = 0; i < idx.size(); ++i) idx[i] = i;
    std::vector<uint64_t> sigs(idx.size() * w64_);
    std::vector<float> norms(idx.size());
    if (!idx.empty()) api_->check(api_->get_rows_sig(h_, idx.size(), idx.data(), sigs.data(), norms.data()));
    for (size_t i = 0; i < keys_.size(); ++i) {
      put_str(out, keys_[i]); put_str(out, versions_[i].first); put_u64(out, versions_[i].second);
    }
    out.append(reinterpret_cast<const char*>(sigs.data()), sigs.size() * 8);
    out.append(reinterpret_cast<const char*>(norms.data()), norms.size() * 4);
    return out;
  }
  void unpack(const std::string& blob) {
    size_t p = 0;
    if (get_u64(blob, p) != 0x314E4E4142554Aull) throw jubatus_exception("unpack: not a jubann table blob");
    if (get_u64(blob, p) != hash_num_) throw jubatus_exception("unpack: hash_num differs from the configured engine");
    const uint64_t clock = get_u64(blob, p), n = get_u64(blob, p);
    std::vector<std::string> keys(n);
    std::vector<version_t> versions(n);
    for (uint64_t i = 0; i < n; ++i) {
      keys[i] = get_str(blob, p); versions[i].first = get_str(blob, p); versions[i].second = get_u64(blob, p);
    }
    if (blob.size() < p + n * w64_ * 8 + n * 4) throw jubatus_exception("unpack: truncated blob");
    clear();
    if (n) {
      std::vector<uint64_t> idx(n);
      for (uint64_t i = 0; i < n; ++i) idx[i] = i;
      std::vector<uint64_t> sigs(n * w64_);
      std::vector<float> norms(n);
      memcpy(sigs.data(), blob.data() + p, n * w64_ * 8);
      memcpy(norms.data(), blob.data() + p + n * w64_ * 8, n * 4);
      api_->check(api_->set_rows_sig(h_, n, idx.data(), sigs.data(), norms.data()));
    }
    keys_.swap(keys); versions_.swap(versions); clock_ = clock;
    for (uint64_t i = 0; i < n; ++i) { index_[keys_[i]] = i; dirty_.insert(i); }
  }

 private:
  static void put_u64(std::string& o, uint64_t v) { o.append(reinterpret_cast<const char*>(&v), 8); }
  static void put_str(std::string& o, const std::string& s) { put_u64(o, s.size()); o.append(s); }
  static uint64_t get_u64(const std::string& b, size_t& p) {
    if (b.size() < p + 8) throw jubatus_exception("unpack: truncated blob");
    uint64_t v; memcpy(&v, b.data() + p, 8); p += 8; return v;
  }
  static std::string get_str(const std::string& b, size_t& p) {
    const uint64_t n = get_u64(b, p);
    if (b.size() < p + n) throw jubatus_exception("unpack: truncated blob");
    std::string s(b.data() + p, n); p += n; return s;
  }
  std::shared_ptr<jubann_api> api_;
  jubann_t* h_;
  uint32_t hash_num_, w64_;
  uint64_t clock_;
  std::vector<std::string> keys_;
  std::vector<version_t> versions_;
  std::unordered_map<std::string, uint64_t> index_;
  std::unordered_set<uint64_t> dirty_;
  bool multi_;
};

// ---------------------------------------------------------------------------------------
// mixable_versioned_table
// ---------------------------------------------------------------------------------------
class mixable_versioned_table {
 public:
  explicit mixable_versioned_table(device_table* t) : table_(t) {}
  void update_version(const version_t& v) {
    version_clock::iterator it = vc_.find(v.first);
    if (it == vc_.end()) vc_[v.first] = v.second;
    else if (it->second < v.second) it->second = v.second;
  }
  version_clock get_argument() {
    for (uint64_t i = 0; i < table_->size(); ++i) update_version(table_->get_version(i));
    return vc_;
  }
  void pull(const version_clock& arg, std::vector<packed_row>& diff) const {
    std::vector<uint64_t> idx;
    for (uint64_t i = 0; i < table_->size(); ++i) {
      const version_t& v = table_->get_version(i);
      version_clock::const_iterator it = arg.find(v.first);
      if (it == arg.end() || it->second < v.second) idx.push_back(i);
    }
    table_->get_rows(idx, diff);
  }
  void push(const std::vector<packed_row>& diff) {
    std::vector<version_t> applied;
    table_->set_rows(diff, &applied);
    for (size_t i = 0; i < applied.size(); ++i) update_version(applied[i]);
  }
  // linear_mixable call shape
  void get_diff(std::vector<packed_row>& diff) const { pull(vc_, diff); }
  static void mix(const std::vector<packed_row>& lhs, std::vector<packed_row>& acc) { acc.insert(acc.end(), lhs.begin(), lhs.end()); }
  bool put_diff(const std::vector<packed_row>& diff) { push(diff); return true; }

  // ---- one MIX round between replicas as collectives under the C-ABI -------------------------
  // (linear_mixer.cpp:422-544: get_diff from every server, fold in server order, put_diff on
  // every server -- here one all-gather of every replica's diff, which every replica then applies
  // in rank order.)  The rows' signatures never visit the host: the library packs them on the
  // device and keeps the gathered records there; the host ships only keys and versions (`blob`)
  // and decides which gathered records are newer than what it holds.  `extra` is appended to the
  // blob and handed back per rank (the driver's second mixable: the converter's weight counts).
  struct mix_stats { uint64_t rows_sent, rows_gathered, rows_stored, blob_bytes; };
  mix_stats mix_allgather(const std::string& extra, std::vector<std::string>* extras) {
    const jubann_api& a = table_->api();
    uint32_t rank = 0, world = 1;
    a.check(a.comm_info(table_->handle(), &rank, &world));
    // get_diff: rows newer than the clocks agreed at the last round
    std::vector<uint64_t> idx;
    std::string blob;
    const std::vector<uint64_t> touched = table_->dirty_rows();   // not the whole table: a round costs what changed
    for (size_t t = 0; t < touched.size(); ++t) {
      const uint64_t i = touched[t];
      const version_t& v = table_->get_version(i);
      version_clock::const_iterator it = vc_.find(v.first);
      if (it == vc_.end() || it->second < v.second) {
        idx.push_back(i);
        put_str(blob, table_->get_key(i)); put_str(blob, v.first); put_u64(blob, v.second);
      }
    }
    const uint64_t meta_bytes = blob.size();
    blob += extra;
    std::vector<uint64_t> rows(world), sizes(world);
    std::vector<const void*> blobs(world);
    std::string head;
    put_u64(head, meta_bytes);
    blob.insert(0, head);
    a.check(a.mix_allgather(table_->handle(), idx.size(), idx.data(), blob.data(), blob.size(), rows.data(), sizes.data(), blobs.data()));
    // put_diff: every gathered record in rank order, own ones included (a row two replicas
    // changed ends up the same everywhere: the record applied last wins among equal clocks)
    std::vector<uint32_t> src_rank, src_pos;
    std::vector<uint64_t> dst;
    mix_stats st = {idx.size(), 0, 0, blob.size()};
    if (extras) extras->assign(world, std::string());
    for (uint32_t r = 0; r < world; ++r) {
      if (!sizes[r]) continue;
      const std::string b(static_cast<const char*>(blobs[r]), sizes[r]);
      size_t p = 0;
      const uint64_t mb = get_u64(b, p);
      if (b.size() < 8 + mb) throw jubatus_exception("mix: truncated blob");
      for (uint64_t j = 0; j < rows[r]; ++j) {
        const std::string key = get_str(b, p), owner = get_str(b, p);
        const uint64_t clock = get_u64(b, p);
        const version_t v(owner, clock);
        uint64_t at;
        if (table_->admit(key, v, &at)) { src_rank.push_back(r); src_pos.push_back(static_cast<uint32_t>(j)); dst.push_back(at); }
        update_version(v);
        ++st.rows_gathered;
      }
      if (extras) (*extras)[r] = b.substr(8 + mb);
    }
    st.rows_stored = dst.size();
    if (!dst.empty()) a.check(a.mix_apply(table_->handle(), dst.size(), src_rank.data(), src_pos.data(), dst.data()));
    table_->clear_dirty();   // every row sent is now covered by the agreed clocks
    return st;
  }

 private:
  static void put_u64(std::string& o, uint64_t v) { o.append(reinterpret_cast<const char*>(&v), 8); }
  static void put_str(std::string& o, const std::string& s) { put_u64(o, s.size()); o.append(s); }
  static uint64_t get_u64(const std::string& b, size_t& p) {
    if (b.size() < p + 8) throw jubatus_exception("mix: truncated blob");
    uint64_t v; memcpy(&v, b.data() + p, 8); p += 8; return v;
  }
  static std::string get_str(const std::string& b, size_t& p) {
    const uint64_t n = get_u64(b, p);
    if (b.size() < p + n) throw jubatus_exception("mix: truncated blob");
    std::string s(b.data() + p, n); p += n; return s;
  }
  device_table* table_;
  version_clock vc_;
};

// ---------------------------------------------------------------------------------------
// nearest_neighbor_base and the GPU engine
// ---------------------------------------------------------------------------------------
typedef std::vector<std::pair<std::string, float> > neighbor_result;

class nearest_neighbor_base {
 public:
  virtual ~nearest_neighbor_base() {}
  virtual std::string type() const = 0;
  virtual void clear() = 0;
  virtual void set_row(const std::string& id, const sfv_t& sfv) = 0;
  virtual void neighbor_row(const sfv_t& query, neighbor_result& ids, uint64_t ret_num) const = 0;
  virtual void neighbor_row(const std::string& query_id, neighbor_result& ids, uint64_t ret_num) const = 0;
  virtual float calc_similarity(float distance) const { return 1 - distance; }
  void similar_row(const sfv_t& query, neighbor_result& ids, uint64_t ret_num) const {
    neighbor_row(query, ids, ret_num);
    for (size_t i = 0; i < ids.size(); ++i) ids[i].second = calc_similarity(ids[i].second);
  }
  void similar_row(const std::string& query_id, neighbor_result& ids, uint64_t ret_num) const {
    neighbor_row(query_id, ids, ret_num);
    for (size_t i = 0; i < ids.size(); ++i) ids[i].second = calc_similarity(ids[i].second);
  }
  virtual void get_all_row_ids(std::vector<std::string>& ids) const = 0;
};

struct csr_batch {
  std::vector<uint64_t> indptr;
  std::vector<uint32_t> seed32;
  std::vector<uint64_t> keyhash64;
  std::vector<float> vals;
  csr_batch() : indptr(1, 0) {}
  void add(const sfv_t& sfv) {
    for (size_t i = 0; i < sfv.size(); ++i) {
      const uint64_t k = calc_string_hash(sfv[i].first);
      keyhash64.push_back(k);
      seed32.push_back(static_cast<uint32_t>(k));
      vals.push_back(sfv[i].second);
    }
    indptr.push_back(vals.size());
  }
  uint64_t rows() const { return indptr.size() - 1; }
};

class gpu_nearest_neighbor : public nearest_neighbor_base {
 public:
  gpu_nearest_neighbor(const std::string& method, uint32_t hash_num, const std::shared_ptr<device_table>& table,
                       const std::string& my_id)
      : method_(method), hash_num_(hash_num), table_(table), my_id_(my_id), mixable_(table.get()) {}

  std::string type() const { return method_; }
  void clear() { std::unique_lock<std::shared_mutex> lk(rw_); table_->clear(); }
  std::shared_ptr<device_table> get_table() { return table_; }
  std::shared_ptr<const device_table> get_const_table() const { return table_; }
  mixable_versioned_table* get_mixable() { return &mixable_; }
  // MIX round with the table's write lock held (push_mixer.cpp:458 takes the model's wlock)
  mixable_versioned_table::mix_stats mix_allgather(const std::string& extra, std::vector<std::string>* extras) {
    std::unique_lock<std::shared_mutex> lk(rw_);
    return mixable_.mix_allgather(extra, extras);
  }
  float calc_similarity(float d) const { return method_ == "euclid_lsh" ? -d : 1 - d; }

  void set_row(const std::string& id, const sfv_t& sfv) {
    std::vector<std::string> ids(1, id);
    std::vector<sfv_t> rows(1, sfv);
    set_rows(ids, rows);
  }
  // batched set_row: one device call for the whole batch (same result as calling set_row in order)
  void set_rows(const std::vector<std::string>& ids, const std::vector<sfv_t>& sfvs) {
    if (ids.empty()) return;
    csr_batch b;
    for (size_t i = 0; i < sfvs.size(); ++i) b.add(sfvs[i]);
    // like column_table's own rw-mutex: writers exclusive, the lock-free RPC readers shared
    std::unique_lock<std::shared_mutex> lk(rw_);
    device_table::assign_journal undo;
    std::vector<uint64_t> idx = table_->assign(ids, my_id_, &undo);
    const jubann_api& a = table_->api();
    // a repeated id repeats its index: the library stores the last entry (column_table::add in order)
    const int rc = a.set_rows_csr(table_->handle(), idx.size(), idx.data(), b.indptr.data(), b.seed32.data(),
                                  b.keyhash64.data(), b.vals.data());
    if (rc) {   // the device table is as it was: so are keys, versions and clock
      table_->rollback(undo);
      a.check(rc);
    }
  }
  // set_row(id, sfv) followed by neighbor_row(id, ret_num) -- what light_lof::set_row does back to
  // back -- as one device call when the table lives on one GPU (jubann_set_row_query_csr); the same
  // answers as the two calls, which it falls back to otherwise.
  void set_row_and_neighbors(const std::string& id, const sfv_t& sfv, neighbor_result& out, uint64_t ret_num) {
    const jubann_api& a = table_->api();
    if (!a.set_row_query_csr || table_->multi_device()) {
      set_row(id, sfv);
      neighbor_row(id, out, ret_num);
      return;
    }
    csr_batch b;
    b.add(sfv);
    std::unique_lock<std::shared_mutex> lk(rw_);
    device_table::assign_journal undo;
    std::vector<std::string> ids(1, id);
    const std::vector<uint64_t> idx = table_->assign(ids, my_id_, &undo);
    const uint32_t k = clamp_k(ret_num);
    std::vector<uint64_t> oi(std::max<uint32_t>(k, 1));
    std::vector<float> os(oi.size());
    std::vector<uint32_t> oc(1);
    const int rc = a.set_row_query_csr(table_->handle(), idx[0], b.vals.size(), b.seed32.data(), b.keyhash64.data(), b.vals.data(), k, 0,
                                       oi.data(), os.data(), oc.data());
    if (rc) {
      table_->rollback(undo);
      a.check(rc);
    }
    std::vector<neighbor_result> res;
    emit(1, k, oi, os, oc, std::vector<bool>(), res);
    out.swap(res[0]);
  }
  void neighbor_row(const sfv_t& query, neighbor_result& ids, uint64_t ret_num) const {
    std::vector<sfv_t> q(1, query);
    std::vector<neighbor_result> out;
    neighbor_rows(q, out, ret_num, false);
    ids.swap(out[0]);
  }
  void neighbor_row(const std::string& query_id, neighbor_result& ids, uint64_t ret_num) const {
    std::vector<std::string> q(1, query_id);
    std::vector<neighbor_result> out;
    neighbor_rows_from_ids(q, out, ret_num, false);
    ids.swap(out[0]);
  }
  // batched queries (what the server-side micro-batcher would call)
  void neighbor_rows(const std::vector<sfv_t>& queries, std::vector<neighbor_result>& out, uint64_t ret_num,
                     bool similar) const {
    csr_batch b;
    for (size_t i = 0; i < queries.size(); ++i) b.add(queries[i]);
    std::shared_lock<std::shared_mutex> lk(rw_);
    const uint32_t k = clamp_k(ret_num);
    std::vector<uint64_t> oi(queries.size() * std::max<uint32_t>(k, 1));
    std::vector<float> os(oi.size());
    std::vector<uint32_t> oc(queries.size());
    const jubann_api& a = table_->api();
    a.check(a.query_csr(table_->handle(), queries.size(), b.indptr.data(), b.seed32.data(), b.keyhash64.data(),
                        b.vals.data(), k, similar ? JUBANN_SIMILAR : 0, oi.data(), os.data(), oc.data()));
    emit(queries.size(), k, oi, os, oc, std::vector<bool>(), out);
  }
  // unknown ids give an empty result, not an error
  void neighbor_rows_from_ids(const std::vector<std::string>& query_ids, std::vector<neighbor_result>& out,
                              uint64_t ret_num, bool similar) const {
    std::shared_lock<std::shared_mutex> lk(rw_);
    std::vector<uint64_t> rows;
    std::vector<bool> known(query_ids.size());
    for (size_t i = 0; i < query_ids.size(); ++i) {
      std::pair<bool, uint64_t> hit = table_->exact_match(query_ids[i]);
      known[i] = hit.first;
      if (hit.first) rows.push_back(hit.second);
    }
    const uint32_t k = clamp_k(ret_num);
    std::vector<uint64_t> oi(rows.size() * std::max<uint32_t>(k, 1));
    std::vector<float> os(oi.size());
    std::vector<uint32_t> oc(rows.size());
    if (!rows.empty()) {
      const jubann_api& a = table_->api();
      a.check(a.query_rows(table_->handle(), rows.size(), rows.data(), k, similar ? JUBANN_SIMILAR : 0, oi.data(),
                           os.data(), oc.data()));
    }
    emit(query_ids.size(), k, oi, os, oc, known, out);
  }
  // ---- the same queries with row indices instead of ids in and out (no strings built): what an
  // embedder that keeps its own per-row columns needs (light_lof's kdist / lrd) ----
  struct raw_neighbors {
    uint32_t k;                       // row stride of idx / dist
    std::vector<uint64_t> idx;        // [n][k] row indices
    std::vector<float> dist;          // [n][k]
    std::vector<uint32_t> count;      // [n]
    void reset(size_t n, uint32_t kk) { k = std::max<uint32_t>(kk, 1); idx.assign(n * k, 0); dist.assign(n * k, 0.0f); count.assign(n, 0); }
  };
  void neighbor_rows_raw(const std::vector<uint64_t>& rows, uint64_t ret_num, raw_neighbors& out) const {
    std::shared_lock<std::shared_mutex> lk(rw_);
    const uint32_t k = clamp_k(ret_num);
    out.reset(rows.size(), k);
    if (rows.empty()) return;
    const jubann_api& a = table_->api();
    a.check(a.query_rows(table_->handle(), rows.size(), rows.data(), k, 0, out.idx.data(), out.dist.data(), out.count.data()));
  }
  void neighbor_row_raw(const sfv_t& query, uint64_t ret_num, raw_neighbors& out) const {
    csr_batch b;
    b.add(query);
    std::shared_lock<std::shared_mutex> lk(rw_);
    const uint32_t k = clamp_k(ret_num);
    out.reset(1, k);
    const jubann_api& a = table_->api();
    a.check(a.query_csr(table_->handle(), 1, b.indptr.data(), b.seed32.data(), b.keyhash64.data(), b.vals.data(), k, 0,
                        out.idx.data(), out.dist.data(), out.count.data()));
  }
  // set_row(id, sfv) then neighbor_row(id): *row = the row's index (see set_row_and_neighbors)
  void set_row_and_neighbors_raw(const std::string& id, const sfv_t& sfv, uint64_t ret_num, uint64_t* row, raw_neighbors& out) {
    const jubann_api& a = table_->api();
    if (!a.set_row_query_csr || table_->multi_device()) {
      set_row(id, sfv);
      *row = table_->exact_match(id).second;
      neighbor_rows_raw(std::vector<uint64_t>(1, *row), ret_num, out);
      return;
    }
    csr_batch b;
    b.add(sfv);
    std::unique_lock<std::shared_mutex> lk(rw_);
    device_table::assign_journal undo;
    std::vector<std::string> ids(1, id);
    const std::vector<uint64_t> idx = table_->assign(ids, my_id_, &undo);
    *row = idx[0];
    const uint32_t k = clamp_k(ret_num);
    out.reset(1, k);
    const int rc = a.set_row_query_csr(table_->handle(), idx[0], b.vals.size(), b.seed32.data(), b.keyhash64.data(), b.vals.data(), k, 0,
                                       out.idx.data(), out.dist.data(), out.count.data());
    if (rc) {
      table_->rollback(undo);
      a.check(rc);
    }
  }
  void get_all_row_ids(std::vector<std::string>& ids) const { std::shared_lock<std::shared_mutex> lk(rw_); ids = table_->get_all_row_ids(); }
  // nearest_neighbor_base::pack / unpack delegate to the table
  std::string pack() const { std::shared_lock<std::shared_mutex> lk(rw_); return table_->pack(); }
  void unpack(const std::string& blob) { std::unique_lock<std::shared_mutex> lk(rw_); table_->unpack(blob); }
  bool delete_row(const std::string& id) { std::unique_lock<std::shared_mutex> lk(rw_); return table_->delete_row(id); }

 private:
  uint32_t clamp_k(uint64_t ret_num) const {
    // ret_num > rows returns all rows; no result list is longer than the table
    return static_cast<uint32_t>(std::min<uint64_t>(ret_num, std::max<uint64_t>(table_->size(), 1)));
  }
  void emit(size_t nq, uint32_t k, const std::vector<uint64_t>& oi, const std::vector<float>& os,
            const std::vector<uint32_t>& oc, const std::vector<bool>& known, std::vector<neighbor_result>& out) const {
    out.assign(nq, neighbor_result());
    size_t r = 0;
    for (size_t q = 0; q < nq; ++q) {
      if (!known.empty() && !known[q]) continue;
      for (uint32_t j = 0; j < oc[r]; ++j)
        out[q].push_back(std::make_pair(table_->get_key(oi[r * std::max<uint32_t>(k, 1) + j]), os[r * std::max<uint32_t>(k, 1) + j]));
      ++r;
    }
  }
  std::string method_;
  uint32_t hash_num_;
  std::shared_ptr<device_table> table_;
  std::string my_id_;
  mixable_versioned_table mixable_;
  mutable std::shared_mutex rw_;
};

// {"hash_num": int [, "threads": int]}; unknown keys are rejected (config_cast_check)
struct nn_parameter {
  int32_t hash_num;
  int32_t threads;
  nn_parameter() : hash_num(64), threads(1) {}
  static nn_parameter from_json(const json::ptr& p) {
    nn_parameter c;
    if (!p) return c;
    if (p->kind != json::value::OBJ) throw config_exception("parameter must be an object");
    for (size_t i = 0; i < p->obj.size(); ++i) {
      const std::string& k = p->obj[i].first;
      const json::ptr& v = p->obj[i].second;
      if (k == "hash_num" || k == "threads") {
        if (v->kind != json::value::NUM || !v->is_integer) throw config_exception("\"" + k + "\" must be an integer");
        (k == "hash_num" ? c.hash_num : c.threads) = static_cast<int32_t>(v->num);
      } else {
        throw config_exception("unknown parameter key: " + k);
      }
    }
    return c;
  }
};

// core::nearest_neighbor::create_nearest_neighbor(method, param, table, my_id).  The device table
// is created here because its geometry depends on (method, hash_num); `threads` is accepted and
// has no effect (the GPU scan is always parallel).
// The GPUs a server process uses: the list given by the embedding code, else $JUBANN_DEVICES
// ("0,1,2,3"), else the one device asked for.
inline std::vector<int> devices_from_env(int device) {
  std::vector<int> out;
  const char* e = ::getenv("JUBANN_DEVICES");
  if (e && *e) {
    const std::string s(e);
    size_t p = 0;
    while (p < s.size()) {
      size_t q = s.find(',', p);
      if (q == std::string::npos) q = s.size();
      if (q > p) out.push_back(atoi(s.substr(p, q - p).c_str()));
      p = q + 1;
    }
  }
  if (out.empty()) out.push_back(device);
  return out;
}

inline std::shared_ptr<gpu_nearest_neighbor> create_nearest_neighbor(const std::string& method, const json::ptr& param,
                                                                     const std::shared_ptr<jubann_api>& api,
                                                                     const std::string& my_id, const std::vector<int>& devices) {
  if (method != "lsh" && method != "euclid_lsh" && method != "minhash") throw unsupported_method(method);
  nn_parameter p = nn_parameter::from_json(param);
  if (p.hash_num < 1) throw invalid_parameter("hash_num should be >= 1");
  std::shared_ptr<device_table> table(new device_table(api, method, static_cast<uint32_t>(p.hash_num), devices));
  return std::shared_ptr<gpu_nearest_neighbor>(new gpu_nearest_neighbor(method, p.hash_num, table, my_id));
}
inline std::shared_ptr<gpu_nearest_neighbor> create_nearest_neighbor(const std::string& method, const json::ptr& param,
                                                                     const std::shared_ptr<jubann_api>& api,
                                                                     const std::string& my_id, int device = 0) {
  return create_nearest_neighbor(method, param, api, my_id, std::vector<int>(1, device));
}

// ---------------------------------------------------------------------------------------
// driver::nearest_neighbor and the server facade
// ---------------------------------------------------------------------------------------
namespace driver {
class nearest_neighbor {
 public:
  nearest_neighbor(const std::shared_ptr<gpu_nearest_neighbor>& nn, const std::shared_ptr<datum_to_fv_converter>& conv)
      : nn_(nn), converter_(conv) {}
  std::shared_ptr<device_table> get_table() { return nn_->get_table(); }
  gpu_nearest_neighbor* engine() { return nn_.get(); }
  void set_row(const std::string& id, const datum& d) { sfv_t v; converter_->convert_and_update_weight(d, v); nn_->set_row(id, v); }
  void set_rows(const std::vector<std::string>& ids, const std::vector<datum>& ds) {
    std::vector<sfv_t> v(ds.size());
    for (size_t i = 0; i < ds.size(); ++i) converter_->convert_and_update_weight(ds[i], v[i]);
    nn_->set_rows(ids, v);
  }
  neighbor_result neighbor_row_from_id(const std::string& id, size_t size) { neighbor_result r; nn_->neighbor_row(id, r, size); return r; }
  neighbor_result neighbor_row_from_datum(const datum& d, size_t size) { sfv_t v; converter_->convert(d, v); neighbor_result r; nn_->neighbor_row(v, r, size); return r; }
  neighbor_result similar_row(const std::string& id, size_t n) { neighbor_result r; nn_->similar_row(id, r, n); return r; }
  neighbor_result similar_row(const datum& d, size_t n) { sfv_t v; converter_->convert(d, v); neighbor_result r; nn_->similar_row(v, r, n); return r; }
  std::vector<neighbor_result> neighbor_rows_from_datums(const std::vector<datum>& ds, size_t size, bool similar) {
    std::vector<sfv_t> v(ds.size());
    for (size_t i = 0; i < ds.size(); ++i) converter_->convert(ds[i], v[i]);
    std::vector<neighbor_result> out;
    nn_->neighbor_rows(v, out, size, similar);
    return out;
  }
  std::vector<std::string> get_all_rows() { std::vector<std::string> r; nn_->get_all_row_ids(r); return r; }
  void clear() { nn_->clear(); converter_->clear_weights(); }
  mixable_versioned_table* get_mixable() { return nn_->get_mixable(); }
  datum_to_fv_converter& converter_mut() { return *converter_; }
  // One MIX round of this driver's two mixables (the versioned table, the converter's weight
  // counts) with the other replicas of the communicator the table's handle has joined: one
  // collective exchange under the C-ABI (see mixable_versioned_table::mix_allgather).
  mixable_versioned_table::mix_stats mix_allgather() {
    weight_manager::counts mine;
    converter_->weights().get_diff(mine);
    std::vector<std::string> extras;
    mixable_versioned_table::mix_stats st = nn_->mix_allgather(weight_manager::pack_counts(mine), &extras);
    weight_manager::counts acc;
    for (size_t r = 0; r < extras.size(); ++r)
      if (!extras[r].empty()) weight_manager::mix(weight_manager::unpack_counts(extras[r]), acc);
    converter_->weights().put_diff(acc);
    return st;
  }
  // driver->pack / unpack (save_load.cpp:129,280): the table, then the converter's global weights
  std::string pack() const {
    const std::string t = nn_->pack(), w = converter_->weights().pack();
    std::string b("JUBANND1");
    uint64_t n = t.size();
    char len[8]; memcpy(len, &n, 8);
    b.append(len, 8); b += t; b += w;
    return b;
  }
  void unpack(const std::string& blob) {
    if (blob.size() < 16 || blob.compare(0, 8, "JUBANND1") != 0) throw jubatus_exception("unpack: not a jubann driver blob");
    uint64_t n; memcpy(&n, blob.data() + 8, 8);
    if (blob.size() < 16 + n) throw jubatus_exception("unpack: truncated blob");
    nn_->unpack(blob.substr(16, n));
    converter_->weights().unpack(blob.substr(16 + n));
  }
  const datum_to_fv_converter& converter() const { return *converter_; }
 private:
  std::shared_ptr<gpu_nearest_neighbor> nn_;
  std::shared_ptr<datum_to_fv_converter> converter_;
};
}  // namespace driver

// ---------------------------------------------------------------------------------------
// query_batcher: turns the server's call pattern -- N RPC worker threads each asking for one
// query without the model lock (NOLOCK_, nearest_neighbor_impl.cpp:75-102; --thread,
// server_util.cpp:155) -- into the batched launches the kernels are measured at.
// The first thread to arrive leads: it waits until `max_batch` requests have queued or `max_wait`
// has passed, takes the whole queue, runs ONE neighbor_rows call per distinct (size, similar)
// and hands every follower its own result.
// ---------------------------------------------------------------------------------------
class query_batcher {
 public:
  query_batcher(gpu_nearest_neighbor* nn, size_t max_batch, std::chrono::microseconds max_wait)
      : nn_(nn), max_batch_(max_batch < 1 ? 1 : max_batch), max_wait_(max_wait), leader_(false), batches_(0), queries_(0) {}

  neighbor_result query(const sfv_t& q, uint64_t ret_num, bool similar) {
    request r;
    r.q = &q; r.ret_num = ret_num; r.similar = similar; r.done = false;
    std::unique_lock<std::mutex> lk(mu_);
    queue_.push_back(&r);
    if (queue_.size() >= max_batch_) cv_.notify_all();
    while (!r.done) {
      if (!leader_) {
        leader_ = true;   // lead one batch
        const std::chrono::steady_clock::time_point deadline = std::chrono::steady_clock::now() + max_wait_;
        while (queue_.size() < max_batch_ && cv_.wait_until(lk, deadline) != std::cv_status::timeout) {}
        std::vector<request*> batch;
        batch.swap(queue_);
        lk.unlock();
        run(batch);
        lk.lock();
        for (size_t i = 0; i < batch.size(); ++i) batch[i]->done = true;
        ++batches_; queries_ += batch.size();
        leader_ = false;
        cv_.notify_all();
      } else {
        cv_.wait(lk);
      }
    }
    if (!r.error.empty()) throw jubatus_exception(r.error);
    return r.result;
  }
  uint64_t batches() const { return batches_; }
  uint64_t queries() const { return queries_; }

 private:
  struct request {
    const sfv_t* q; uint64_t ret_num; bool similar; bool done;
    neighbor_result result; std::string error;
  };
  void run(const std::vector<request*>& batch) {
    std::vector<bool> taken(batch.size(), false);
    for (size_t i = 0; i < batch.size(); ++i) {
      if (taken[i]) continue;
      std::vector<size_t> members;
      std::vector<sfv_t> qs;
      for (size_t j = i; j < batch.size(); ++j) {
        if (!taken[j] && batch[j]->ret_num == batch[i]->ret_num && batch[j]->similar == batch[i]->similar) {
          taken[j] = true; members.push_back(j); qs.push_back(*batch[j]->q);
        }
      }
      try {
        std::vector<neighbor_result> out;
        nn_->neighbor_rows(qs, out, batch[i]->ret_num, batch[i]->similar);
        for (size_t m = 0; m < members.size(); ++m) batch[members[m]]->result.swap(out[m]);
      } catch (const std::exception& e) {
        for (size_t m = 0; m < members.size(); ++m) batch[members[m]]->error = e.what();
      }
    }
  }
  gpu_nearest_neighbor* nn_;
  size_t max_batch_;
  std::chrono::microseconds max_wait_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::vector<request*> queue_;
  bool leader_;
  uint64_t batches_, queries_;
};

// jubatus::server::nearest_neighbor_serv: config -> engine, RPC-named pass-throughs
class nearest_neighbor_serv {
 public:
  explicit nearest_neighbor_serv(const std::shared_ptr<jubann_api>& api, const std::string& my_id = "", int device = 0)
      : api_(api), my_id_(my_id), devices_(1, device), update_row_cnt_(0) {}
  // a server process that owns several GPUs: the table is sharded over them inside the library
  nearest_neighbor_serv(const std::shared_ptr<jubann_api>& api, const std::string& my_id, const std::vector<int>& devices)
      : api_(api), my_id_(my_id), devices_(devices), update_row_cnt_(0) {}
  void set_config(const std::string& config) {
    json::ptr root;
    try { root = json::parse(config); } catch (const json::parse_error& e) { throw config_exception(e.what()); }
    if (root->kind != json::value::OBJ) throw config_exception("config must be an object");
    for (size_t i = 0; i < root->obj.size(); ++i) {
      const std::string& k = root->obj[i].first;
      if (k != "method" && k != "parameter" && k != "converter") throw config_exception("unknown config key: " + k);
    }
    json::ptr m = root->find("method");
    if (!m || m->kind != json::value::STR) throw config_exception("\"method\" is required");
    json::ptr conv = root->find("converter");
    if (!conv) throw config_exception("\"converter\" is required");
    std::shared_ptr<datum_to_fv_converter> c = datum_to_fv_converter::from_config(conv);
    std::shared_ptr<gpu_nearest_neighbor> nn = create_nearest_neighbor(m->str, root->find("parameter"), api_, my_id_, devices_);
    nearest_neighbor_.reset(new driver::nearest_neighbor(nn, c));
    config_ = config;
  }
  std::string get_config() const { check_set_config(); return config_; }
  bool clear() { check_set_config(); update_row_cnt_ = 0; nearest_neighbor_->clear(); return true; }
  bool set_row(const std::string& id, const datum& d) { check_set_config(); ++update_row_cnt_; nearest_neighbor_->set_row(id, d); return true; }
  neighbor_result neighbor_row_from_id(const std::string& id, size_t size) { check_set_config(); return nearest_neighbor_->neighbor_row_from_id(id, size); }
  neighbor_result neighbor_row_from_datum(const datum& d, size_t size) {
    check_set_config();
    if (batcher_) { sfv_t v; nearest_neighbor_->converter().convert(d, v); return batcher_->query(v, size, false); }
    return nearest_neighbor_->neighbor_row_from_datum(d, size);
  }
  neighbor_result similar_row_from_id(const std::string& id, size_t n) { check_set_config(); return nearest_neighbor_->similar_row(id, n); }
  neighbor_result similar_row_from_datum(const datum& d, size_t n) {
    check_set_config();
    if (batcher_) { sfv_t v; nearest_neighbor_->converter().convert(d, v); return batcher_->query(v, n, true); }
    return nearest_neighbor_->similar_row(d, n);
  }
  // coalesce concurrent *_from_datum calls into batched device launches (off by default)
  void enable_batching(size_t max_batch, uint64_t max_wait_us) {
    check_set_config();
    batcher_.reset(new query_batcher(nearest_neighbor_->engine(), max_batch, std::chrono::microseconds(max_wait_us)));
  }
  uint64_t batches_run() const { return batcher_ ? batcher_->batches() : 0; }
  uint64_t queries_batched() const { return batcher_ ? batcher_->queries() : 0; }
  std::vector<std::string> get_all_rows() const { check_set_config(); return nearest_neighbor_->get_all_rows(); }
  uint64_t update_row_cnt() const { return update_row_cnt_; }
  driver::nearest_neighbor* get_driver() { check_set_config(); return nearest_neighbor_.get(); }

  // ---- model files (framework/server_base.cpp:135-212, framework/save_load.cpp) ----
  // where `save(id)` / `load(id)` put their file: <datadir>/<eth>_<port>_<type>_<id>.jubatus
  // (build_local_path, server_base.cpp:41-50; defaults of server_util.cpp: /tmp, port 9199)
  void set_storage(const std::string& datadir, const std::string& eth, int port) { datadir_ = datadir; eth_ = eth; port_ = port; }
  static const char* type() { return "nearest_neighbor"; }
  // nearest_neighbor_serv.cpp:77-79 returns 1 for a payload in jubatus_core's msgpack form.  The payload
  // written here is this build's own table blob, so it carries a version of its own ("JBN", 1): a stock
  // server refuses such a file at the version check ("user data version mismatched"), and this one refuses
  // a stock file the same way, instead of either of them failing somewhere inside unpack.
  uint64_t user_data_version() const { return 0x4A424E01ull; }
  std::string local_path(const std::string& id) const {
    std::ostringstream p; p << datadir_ << '/' << eth_ << '_' << port_ << '_' << type() << '_' << id << ".jubatus";
    return p.str();
  }
  // returns {server identifier "<eth>_<port>" -> path}, like the `save` RPC
  std::map<std::string, std::string> save(const std::string& id) {
    check_set_config();
    if (id.empty()) throw jubatus_exception("empty id is not allowed");
    const std::string path = local_path(id);
    {
      std::ofstream ofs(path.c_str(), std::ios::binary | std::ios::trunc);
      if (!ofs) throw jubatus_exception("cannot open output file: " + path);
      try {
        model_file::save(ofs, type(), id, config_, user_data_version(), nearest_neighbor_->pack());
        ofs.close();
        if (!ofs) throw std::ios_base::failure("close");
      } catch (const std::ios_base::failure&) {
        remove(path.c_str());   // no dirty model file is left behind (server_base.cpp:170-182)
        throw jubatus_exception("cannot write output file: " + path);
      }
    }
    last_saved_path_ = path;
    std::map<std::string, std::string> ret;
    std::ostringstream ident; ident << eth_ << '_' << port_;
    ret[ident.str()] = path;
    return ret;
  }
  // `load` RPC: the file must carry this server's type and (semantically) its config
  bool load(const std::string& id) {
    if (id.empty()) throw jubatus_exception("empty id is not allowed");
    load_file_impl(local_path(id), false);
    return true;
  }
  // --model_file at start-up: the file's config replaces the current one (server_base.cpp:207-209)
  void load_file(const std::string& path) { load_file_impl(path, true); }
  const std::string& last_saved_path() const { return last_saved_path_; }
  const std::string& last_loaded_path() const { return last_loaded_path_; }
 private:
  void load_file_impl(const std::string& path, bool overwrite_config) {
    std::ifstream ifs(path.c_str(), std::ios::binary);
    if (!ifs) throw jubatus_exception("cannot open input file: " + path);
    model_file::system_data sd;
    std::string blob;
    try {
      if (!overwrite_config) check_set_config();
      blob = model_file::load(ifs, type(), config_, overwrite_config, user_data_version(), &sd);
    } catch (const model_file::format_error& e) {
      throw jubatus_exception(e.what());
    } catch (const std::ios_base::failure&) {
      throw jubatus_exception("cannot read input file: " + path);
    }
    if (overwrite_config) set_config(sd.config);
    nearest_neighbor_->unpack(blob);
    last_loaded_path_ = path;
  }
  void check_set_config() const { if (!nearest_neighbor_) throw config_exception("config is not set"); }
  std::shared_ptr<jubann_api> api_;
  std::string my_id_;
  std::vector<int> devices_;
  std::string config_;
  uint64_t update_row_cnt_;
  std::shared_ptr<driver::nearest_neighbor> nearest_neighbor_;
  std::shared_ptr<query_batcher> batcher_;
  std::string datadir_ = "/tmp", eth_ = "127.0.0.1", last_saved_path_, last_loaded_path_;
  int port_ = 9199;
};

// ---------------------------------------------------------------------------------------
// core::unlearner -- keeps the model at most `max_size` rows.  Config surface:
// "unlearner": "lru" | "random", "unlearner_parameter": {"max_size": N [, "seed": S]}
// (config/recommender/nearest_neighbor_recommender_euclid_lsh_unlearn_lru.json:21-24,
// config/anomaly/light_lof_unlearn_lru.json:23-26).  touch(id) marks a row as just used; when a
// new id would exceed max_size the policy picks a victim and the owner's callback deletes it
// (column_table::delete_row: swap with the last row).  [core] unlearner/{lru,random}_unlearner --
// restated from recall: lru evicts the least recently touched id, random a uniformly drawn one.
// ---------------------------------------------------------------------------------------
class unlearner_base {
 public:
  typedef std::function<void(const std::string&)> callback_t;
  virtual ~unlearner_base() {}
  virtual std::string type() const = 0;
  virtual void clear() = 0;
  virtual bool touch(const std::string& id) = 0;      // false: cannot make room (max_size 0 is rejected at construction)
  virtual bool remove(const std::string& id) = 0;
  virtual bool exists_in_memory(const std::string& id) const = 0;
  virtual size_t size() const = 0;
  void set_callback(const callback_t& cb) { callback_ = cb; }
 protected:
  callback_t callback_;
};

class lru_unlearner : public unlearner_base {
 public:
  explicit lru_unlearner(size_t max_size) : max_size_(max_size) {
    if (max_size_ == 0) throw invalid_parameter("max_size must be a positive integer");
  }
  std::string type() const { return "lru_unlearner"; }
  void clear() { lru_.clear(); pos_.clear(); }
  bool touch(const std::string& id) {
    std::unordered_map<std::string, std::list<std::string>::iterator>::iterator it = pos_.find(id);
    if (it != pos_.end()) {
      lru_.splice(lru_.begin(), lru_, it->second);   // most recently used first
      return true;
    }
    if (lru_.size() >= max_size_) {
      const std::string victim = lru_.back();
      pos_.erase(victim);
      lru_.pop_back();
      if (callback_) callback_(victim);
    }
    lru_.push_front(id);
    pos_[id] = lru_.begin();
    return true;
  }
  bool remove(const std::string& id) {
    std::unordered_map<std::string, std::list<std::string>::iterator>::iterator it = pos_.find(id);
    if (it == pos_.end()) return false;
    lru_.erase(it->second);
    pos_.erase(it);
    return true;
  }
  bool exists_in_memory(const std::string& id) const { return pos_.count(id) != 0; }
  size_t size() const { return lru_.size(); }
 private:
  size_t max_size_;
  std::list<std::string> lru_;
  std::unordered_map<std::string, std::list<std::string>::iterator> pos_;
};

class random_unlearner : public unlearner_base {
 public:
  random_unlearner(size_t max_size, uint64_t seed) : max_size_(max_size), state_(seed ? seed : 0x9E3779B97F4A7C15ull) {
    if (max_size_ == 0) throw invalid_parameter("max_size must be a positive integer");
  }
  std::string type() const { return "random_unlearner"; }
  void clear() { ids_.clear(); pos_.clear(); }
  bool touch(const std::string& id) {
    if (pos_.count(id)) return true;
    if (ids_.size() >= max_size_) {
      const size_t v = static_cast<size_t>(next() % ids_.size());
      const std::string victim = ids_[v];
      ids_[v] = ids_.back();
      pos_[ids_[v]] = v;
      ids_.pop_back();
      pos_.erase(victim);
      if (callback_) callback_(victim);
    }
    pos_[id] = ids_.size();
    ids_.push_back(id);
    return true;
  }
  bool remove(const std::string& id) {
    std::unordered_map<std::string, size_t>::iterator it = pos_.find(id);
    if (it == pos_.end()) return false;
    const size_t v = it->second;
    ids_[v] = ids_.back();
    pos_[ids_[v]] = v;
    ids_.pop_back();
    pos_.erase(id);
    return true;
  }
  bool exists_in_memory(const std::string& id) const { return pos_.count(id) != 0; }
  size_t size() const { return ids_.size(); }
 private:
  uint64_t next() {   // splitmix64
    uint64_t z = (state_ += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  size_t max_size_;
  uint64_t state_;
  std::vector<std::string> ids_;
  std::unordered_map<std::string, size_t> pos_;
};

// create_unlearner(name, config) -- core::unlearner::create_unlearner
inline std::shared_ptr<unlearner_base> create_unlearner(const std::string& name, const json::ptr& param) {
  if (!param || param->kind != json::value::OBJ) throw config_exception("\"unlearner_parameter\" is required with \"unlearner\"");
  int64_t max_size = -1, seed = 0;
  for (size_t i = 0; i < param->obj.size(); ++i) {
    const std::string& k = param->obj[i].first;
    const json::ptr& v = param->obj[i].second;
    if (k == "max_size" || k == "seed") {
      if (v->kind != json::value::NUM || !v->is_integer) throw config_exception("\"" + k + "\" must be an integer");
      (k == "max_size" ? max_size : seed) = static_cast<int64_t>(v->num);
    } else if (k != "sticky_pattern") {
      throw config_exception("unknown unlearner_parameter key: " + k);
    } else {
      throw config_exception("sticky_pattern is not supported by this build");
    }
  }
  if (max_size <= 0) throw invalid_parameter("max_size must be a positive integer");
  if (name == "lru") return std::shared_ptr<unlearner_base>(new lru_unlearner(static_cast<size_t>(max_size)));
  if (name == "random") return std::shared_ptr<unlearner_base>(new random_unlearner(static_cast<size_t>(max_size), static_cast<uint64_t>(seed)));
  throw unsupported_method("unlearner: " + name);
}

// ---------------------------------------------------------------------------------------
// recommender::nearest_neighbor_recommender
// ---------------------------------------------------------------------------------------
class nearest_neighbor_recommender {
 public:
  explicit nearest_neighbor_recommender(const std::shared_ptr<gpu_nearest_neighbor>& nn,
                                        const std::shared_ptr<unlearner_base>& unlearner = std::shared_ptr<unlearner_base>())
      : nn_(nn), unlearner_(unlearner) {
    // the victim leaves both the original rows and the engine's table
    if (unlearner_) unlearner_->set_callback([this](const std::string& victim) { orig_.erase(victim); nn_->delete_row(victim); });
  }
  void clear() { orig_.clear(); nn_->clear(); if (unlearner_) unlearner_->clear(); }
  const unlearner_base* get_unlearner() const { return unlearner_.get(); }
  // update_row merges the diff into the stored original row, then re-hashes the FULL row
  void update_row(const std::string& id, const sfv_t& diff) {
    if (unlearner_ && !unlearner_->touch(id)) throw jubatus_exception("cannot add new row as number of sticky rows reached the maximum size of unlearner: " + id);
    std::vector<std::pair<std::string, float> >& row = orig_[id];
    for (size_t i = 0; i < diff.size(); ++i) {
      bool found = false;
      for (size_t j = 0; j < row.size(); ++j) if (row[j].first == diff[i].first) { row[j].second = diff[i].second; found = true; break; }
      if (!found) row.push_back(diff[i]);
    }
    nn_->set_row(id, row);
  }
  void clear_row(const std::string& id) { orig_.erase(id); nn_->delete_row(id); if (unlearner_) unlearner_->remove(id); }
  void similar_row(const sfv_t& q, neighbor_result& ids, size_t n) const { nn_->similar_row(q, ids, n); }
  void neighbor_row(const sfv_t& q, neighbor_result& ids, size_t n) const { nn_->neighbor_row(q, ids, n); }
  // *_from_id re-hashes the stored original row (not the stored signature); unknown id -> empty
  void similar_row(const std::string& id, neighbor_result& ids, size_t n) const {
    std::unordered_map<std::string, sfv_t>::const_iterator it = orig_.find(id);
    if (it == orig_.end()) { ids.clear(); return; }
    nn_->similar_row(it->second, ids, n);
  }
  void neighbor_row(const std::string& id, neighbor_result& ids, size_t n) const {
    std::unordered_map<std::string, sfv_t>::const_iterator it = orig_.find(id);
    if (it == orig_.end()) { ids.clear(); return; }
    nn_->neighbor_row(it->second, ids, n);
  }
  void decode_row(const std::string& id, sfv_t& out) const {
    std::unordered_map<std::string, sfv_t>::const_iterator it = orig_.find(id);
    if (it == orig_.end()) out.clear(); else out = it->second;
  }
  void get_all_row_ids(std::vector<std::string>& ids) const { nn_->get_all_row_ids(ids); }
  std::string type() const { return "nearest_neighbor_recommender:" + nn_->type(); }
 private:
  std::shared_ptr<gpu_nearest_neighbor> nn_;
  std::shared_ptr<unlearner_base> unlearner_;
  std::unordered_map<std::string, sfv_t> orig_;
};

// ---------------------------------------------------------------------------------------
// classifier::nearest_neighbor_classifier -- method "NN" of jubaclassifier
// (config/classifier/nn.json: {"method": "NN", "parameter": {"method": <engine>, "parameter":
// {...}, "nearest_neighbor_num": k, "local_sensitivity": alpha}}; RPC surface classifier.idl:
// train, classify, get_labels, set_label, delete_label, clear).  Restated from recall of
// [core] classifier/nearest_neighbor_classifier.cpp: every training datum becomes a row whose id
// is "<label>_<serial>"; classify asks for the k nearest rows and adds exp(-alpha * distance)
// to the label each row carries; every known label is reported, unseen ones with score 0.
// ---------------------------------------------------------------------------------------
class nearest_neighbor_classifier {
 public:
  typedef std::vector<std::pair<std::string, float> > classify_result;   // (label, score), label order
  nearest_neighbor_classifier(const std::shared_ptr<gpu_nearest_neighbor>& nn, size_t k, float alpha,
                              const std::shared_ptr<unlearner_base>& unlearner = std::shared_ptr<unlearner_base>())
      : nn_(nn), k_(k), alpha_(alpha), serial_(0), unlearner_(unlearner) {
    if (k_ == 0) throw invalid_parameter("nearest_neighbor_num should be >= 1");
    if (!(alpha_ >= 0.0f)) throw invalid_parameter("local_sensitivity should be >= 0");
    if (unlearner_) unlearner_->set_callback([this](const std::string& victim) { forget(victim); });
  }
  void train(const sfv_t& fv, const std::string& label) {
    const std::string id = label + "_" + std::to_string(serial_++);
    if (unlearner_ && !unlearner_->touch(id)) throw jubatus_exception("no more space available to add new ID: " + id);
    labels_[label] += 1;
    nn_->set_row(id, fv);
  }
  void classify_with_scores(const sfv_t& fv, classify_result& scores) const {
    neighbor_result ids;
    nn_->neighbor_row(fv, ids, k_);
    std::map<std::string, float> m;
    for (std::map<std::string, uint64_t>::const_iterator it = labels_.begin(); it != labels_.end(); ++it) m[it->first] = 0.0f;
    for (size_t i = 0; i < ids.size(); ++i) m[label_of(ids[i].first)] += expf(-alpha_ * ids[i].second);
    scores.assign(m.begin(), m.end());
  }
  // one device call for a whole classify RPC (it takes a list of datums)
  void classify_batch(const std::vector<sfv_t>& fvs, std::vector<classify_result>& out) const {
    std::vector<neighbor_result> nb;
    nn_->neighbor_rows(fvs, nb, k_, false);
    out.resize(fvs.size());
    for (size_t r = 0; r < fvs.size(); ++r) {
      std::map<std::string, float> m;
      for (std::map<std::string, uint64_t>::const_iterator it = labels_.begin(); it != labels_.end(); ++it) m[it->first] = 0.0f;
      for (size_t i = 0; i < nb[r].size(); ++i) m[label_of(nb[r][i].first)] += expf(-alpha_ * nb[r][i].second);
      out[r].assign(m.begin(), m.end());
    }
  }
  std::map<std::string, uint64_t> get_labels() const { return labels_; }
  bool set_label(const std::string& label) { return labels_.insert(std::make_pair(label, static_cast<uint64_t>(0))).second; }
  // delete_label drops the label and every row trained with it
  bool delete_label(const std::string& label) {
    if (labels_.erase(label) == 0) return false;
    std::vector<std::string> all;
    nn_->get_all_row_ids(all);
    for (size_t i = 0; i < all.size(); ++i)
      if (label_of(all[i]) == label) { nn_->delete_row(all[i]); if (unlearner_) unlearner_->remove(all[i]); }
    return true;
  }
  void clear() { nn_->clear(); labels_.clear(); serial_ = 0; if (unlearner_) unlearner_->clear(); }
  gpu_nearest_neighbor* engine() { return nn_.get(); }
  static std::string label_of(const std::string& id) { return id.substr(0, id.rfind('_')); }
 private:
  void forget(const std::string& id) {
    nn_->delete_row(id);
    std::map<std::string, uint64_t>::iterator it = labels_.find(label_of(id));
    if (it != labels_.end() && it->second > 0) it->second -= 1;
  }
  std::shared_ptr<gpu_nearest_neighbor> nn_;
  size_t k_;
  float alpha_;
  uint64_t serial_;
  std::shared_ptr<unlearner_base> unlearner_;
  std::map<std::string, uint64_t> labels_;
};

// ---------------------------------------------------------------------------------------
// anomaly::light_lof -- the 1 + reverse_nearest_neighbor_num re-queries of one add() go to the
// device as ONE batched call (they share the streaming passes)
// ---------------------------------------------------------------------------------------
class light_lof {
 public:
  typedef gpu_nearest_neighbor::raw_neighbors raw_neighbors;
  light_lof(const std::shared_ptr<gpu_nearest_neighbor>& nn, int nearest_neighbor_num, int reverse_nearest_neighbor_num,
            const std::shared_ptr<unlearner_base>& unlearner = std::shared_ptr<unlearner_base>())
      : nn_(nn), nn_num_(nearest_neighbor_num), rnn_num_(reverse_nearest_neighbor_num), unlearner_(unlearner) {
    if (nn_num_ < 2) throw invalid_parameter("nearest_neighbor_num should be >= 2");
    if (rnn_num_ < nn_num_) throw invalid_parameter("reverse_nearest_neighbor_num should be >= nearest_neighbor_num");
    // the victim leaves the engine's table and the (kdist, lrd) table; rows that had it as a
    // neighbour keep their parameters until they are next re-scored
    if (unlearner_) unlearner_->set_callback([this](const std::string& victim) { delete_row(victim); });
  }
  void clear() { nn_->clear(); kdist_.clear(); lrd_.clear(); param_dirty_.clear(); if (unlearner_) unlearner_->clear(); }
  const unlearner_base* get_unlearner() const { return unlearner_.get(); }
  // column_table::delete_row on both tables: the engine's last row moves into the hole, so does its
  // (kdist, lrd) entry
  bool delete_row(const std::string& id) {
    const std::pair<bool, uint64_t> hit = nn_->get_table()->exact_match(id);
    if (!hit.first) return false;
    const uint64_t at = hit.second, last = nn_->get_table()->size() - 1;
    sync_rows();
    nn_->delete_row(id);
    kdist_[at] = kdist_[last]; lrd_[at] = lrd_[last];
    kdist_.pop_back(); lrd_.pop_back();
    const bool last_dirty = param_dirty_.erase(last) != 0;
    param_dirty_.erase(at);
    if (last_dirty && at != last) param_dirty_.insert(at);
    return true;
  }

  // light_lof's second mixable ([core] anomaly/light_lof: a versioned table of (kdist, lrd) beside
  // the engine's; anomaly_serv.cpp:134-138 registers both with the mixer).  The diff is every row
  // whose parameters were (re)computed here since the last round; every replica applies the
  // gathered diffs in rank order, so a row two replicas re-scored ends up the same everywhere.
  std::string pack_param_diff() const {
    std::string out;
    const uint64_t n = param_dirty_.size();
    out.append(reinterpret_cast<const char*>(&n), 8);
    std::vector<std::pair<std::string, uint64_t> > keys;
    for (std::unordered_set<uint64_t>::const_iterator it = param_dirty_.begin(); it != param_dirty_.end(); ++it)
      keys.push_back(std::make_pair(nn_->get_table()->get_key(*it), *it));
    std::sort(keys.begin(), keys.end());
    for (size_t i = 0; i < keys.size(); ++i) {
      const uint64_t len = keys[i].first.size();
      const float kd = kdist_[keys[i].second], lr = lrd_[keys[i].second];
      out.append(reinterpret_cast<const char*>(&len), 8);
      out.append(keys[i].first);
      out.append(reinterpret_cast<const char*>(&kd), 4);
      out.append(reinterpret_cast<const char*>(&lr), 4);
    }
    return out;
  }
  // returns the bytes consumed; rows this replica does not hold (unlearned meanwhile) are skipped
  size_t apply_param_diff(const std::string& b, size_t p = 0) {
    if (b.size() < p + 8) throw jubatus_exception("mix: truncated light_lof diff");
    uint64_t n; memcpy(&n, b.data() + p, 8); p += 8;
    sync_rows();
    for (uint64_t i = 0; i < n; ++i) {
      if (b.size() < p + 8) throw jubatus_exception("mix: truncated light_lof diff");
      uint64_t len; memcpy(&len, b.data() + p, 8); p += 8;
      if (b.size() < p + len + 8) throw jubatus_exception("mix: truncated light_lof diff");
      const std::string key(b.data() + p, len); p += len;
      float kd, lr; memcpy(&kd, b.data() + p, 4); memcpy(&lr, b.data() + p + 4, 4); p += 8;
      const std::pair<bool, uint64_t> hit = nn_->get_table()->exact_match(key);
      if (hit.first) { kdist_[hit.second] = kd; lrd_[hit.second] = lr; }
    }
    return p;
  }
  void clear_param_dirty() { param_dirty_.clear(); }
  // order-independent digest of the (kdist, lrd) table (tests, bench: replicas agree after MIX)
  uint64_t param_digest() const {
    uint64_t acc = 0;
    const size_t n = std::min<size_t>(kdist_.size(), nn_->get_table()->size());
    for (size_t i = 0; i < n; ++i) {
      uint32_t a, b;
      memcpy(&a, &kdist_[i], 4); memcpy(&b, &lrd_[i], 4);
      uint64_t h = calc_string_hash(nn_->get_table()->get_key(i)) ^ (static_cast<uint64_t>(a) << 32 | b);
      h ^= h >> 33; h *= 0xff51afd7ed558ccdULL; h ^= h >> 33;
      acc += h;
    }
    return acc;
  }
  static float calculate_lof(float lrd, const std::vector<float>& neighbor_lrds) {
    if (neighbor_lrds.empty()) return lrd == 0 ? 1.0f : std::numeric_limits<float>::infinity();
    float sum = 0.0f;
    for (size_t i = 0; i < neighbor_lrds.size(); ++i) sum += neighbor_lrds[i];
    if (std::isinf(sum) && std::isinf(lrd)) return 1.0f;
    return sum / (neighbor_lrds.size() * lrd);
  }
  float calc_anomaly_score(const sfv_t& query) const {
    raw_neighbors nb;
    nn_->neighbor_row_raw(query, nn_num_, nb);
    return score_from_neighbors(nb, 0);
  }
  float calc_anomaly_score(const std::string& id) const {
    const std::pair<bool, uint64_t> hit = nn_->get_table()->exact_match(id);
    raw_neighbors nb;
    if (hit.first) nn_->neighbor_rows_raw(std::vector<uint64_t>(1, hit.second), nn_num_, nb);
    else nb.reset(1, 1);       // unknown id: no neighbours (neighbor_row(id) gives an empty list)
    return score_from_neighbors(nb, 0);
  }
  bool set_row(const std::string& id, const sfv_t& sfv) { return set_row_impl(id, sfv, 0); }
  // anomaly add: set_row, then the new row's score (anomaly_serv.cpp:157-176).  The row's neighbour
  // list is already at hand from the re-scoring of set_row and the table has not changed since, so
  // the score costs no further device call; the value is calc_anomaly_score(id)'s.
  float set_row_and_score(const std::string& id, const sfv_t& sfv) {
    raw_neighbors own;
    own.reset(1, 1);
    set_row_impl(id, sfv, &own);
    return score_from_neighbors(own, 0);
  }
  // kdist, then lrd of the listed rows from one batched device call (every row's nn_num_ nearest);
  // want_row's own neighbour list is handed back when asked for
  void update_entries(const std::vector<uint64_t>& rows, const uint64_t* want_row = 0, raw_neighbors* want = 0) {
    raw_neighbors nn;
    nn_->neighbor_rows_raw(rows, nn_num_, nn);
    sync_rows();
    for (size_t i = 0; i < rows.size(); ++i) kdist_[rows[i]] = nn.count[i] ? nn.dist[i * nn.k + nn.count[i] - 1] : 0.0f;
    for (size_t i = 0; i < rows.size(); ++i) {
      const size_t length = std::min<size_t>(nn.count[i], nn_num_);
      float sum_reach = 0.0f;
      for (size_t j = 0; j < length; ++j) sum_reach += std::max(nn.dist[i * nn.k + j], kdist_of(nn.idx[i * nn.k + j]));
      lrd_[rows[i]] = sum_reach == 0 ? std::numeric_limits<float>::infinity() : length / sum_reach;
      param_dirty_.insert(rows[i]);
      if (want_row && want && rows[i] == *want_row) {
        want->reset(1, nn.k);
        want->count[0] = nn.count[i];
        for (uint32_t j = 0; j < nn.count[i]; ++j) { want->idx[j] = nn.idx[i * nn.k + j]; want->dist[j] = nn.dist[i * nn.k + j]; }
      }
    }
  }
  // Bulk load for a table that is built in one go (benchmarks, restoring from raw data): rows go
  // in through the batched set_rows, then kdist for every row, then lrd for every row, each in
  // batched device calls.  Unlike a sequence of set_row calls no row keeps a stale parameter.
  void set_rows_bulk(const std::vector<std::string>& ids, const std::vector<sfv_t>& sfvs, size_t batch = 4096) {
    if (unlearner_) throw jubatus_exception("bulk load is not available with an unlearner (rows must arrive one by one)");
    nn_->set_rows(ids, sfvs);
    sync_rows();
    const uint64_t n = nn_->get_table()->size();
    raw_neighbors nn;
    std::vector<uint64_t> part;
    for (int phase = 0; phase < 2; ++phase) {
      for (uint64_t b = 0; b < n; b += batch) {
        part.resize(std::min<uint64_t>(n, b + batch) - b);
        for (size_t i = 0; i < part.size(); ++i) part[i] = b + i;
        nn_->neighbor_rows_raw(part, nn_num_, nn);
        for (size_t i = 0; i < part.size(); ++i) {
          if (phase == 0) {
            kdist_[part[i]] = nn.count[i] ? nn.dist[i * nn.k + nn.count[i] - 1] : 0.0f;
          } else {
            const size_t length = std::min<size_t>(nn.count[i], nn_num_);
            float sum_reach = 0.0f;
            for (size_t j = 0; j < length; ++j) sum_reach += std::max(nn.dist[i * nn.k + j], kdist_[nn.idx[i * nn.k + j]]);
            lrd_[part[i]] = sum_reach == 0 ? std::numeric_limits<float>::infinity() : length / sum_reach;
            param_dirty_.insert(part[i]);
          }
        }
      }
    }
  }
  bool get_row_parameter(const std::string& id, float* kdist, float* lrd) const {
    const std::pair<bool, uint64_t> hit = nn_->get_table()->exact_match(id);
    if (!hit.first || hit.second >= kdist_.size()) return false;
    *kdist = kdist_[hit.second]; *lrd = lrd_[hit.second];
    return true;
  }
  gpu_nearest_neighbor* engine() { return nn_.get(); }
 private:
  // The (kdist, lrd) columns are kept by row index of the engine's table (two flat arrays, like the
  // float columns of the reference's second column_table) -- an add touches some 350 entries, which
  // in string-keyed hash maps of 10^6 entries was a quarter of its time.  Rows the table has gained
  // since (MIX appends) start without parameters (-1) until their owner's diff arrives.
  void sync_rows() {
    const size_t n = nn_->get_table()->size();
    if (kdist_.size() < n) { kdist_.resize(n, -1.0f); lrd_.resize(n, -1.0f); }
  }
  float kdist_of(uint64_t row) const { return row < kdist_.size() ? kdist_[row] : 0.0f; }
  bool set_row_impl(const std::string& id, const sfv_t& sfv, raw_neighbors* own) {
    if (unlearner_ && !unlearner_->touch(id)) throw jubatus_exception("no more space available to add new ID: " + id);
    std::vector<uint64_t> order;
    std::unordered_set<uint64_t> seen;
    const std::pair<bool, uint64_t> before = nn_->get_table()->exact_match(id);
    raw_neighbors nb;
    if (before.first) {   // an update: the neighbours of the row as it was are re-scored too
      nn_->neighbor_rows_raw(std::vector<uint64_t>(1, before.second), rnn_num_, nb);
      for (uint32_t j = 0; j < nb.count[0]; ++j) if (seen.insert(nb.idx[j]).second) order.push_back(nb.idx[j]);
    }
    // set_row, then the row's reverse-neighbour candidates: one device call
    static const bool trace = ::getenv("JUBANN_HOST_TRACE") != 0;
    static double t_set = 0, t_upd = 0; static uint64_t n_calls = 0;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    uint64_t row = 0;
    nn_->set_row_and_neighbors_raw(id, sfv, rnn_num_, &row, nb);
    std::chrono::steady_clock::time_point t1 = std::chrono::steady_clock::now();
    for (uint32_t j = 0; j < nb.count[0]; ++j) if (seen.insert(nb.idx[j]).second) order.push_back(nb.idx[j]);
    sync_rows();
    kdist_[row] = -1.0f; lrd_[row] = -1.0f;
    if (seen.insert(row).second) order.push_back(row);
    update_entries(order, own ? &row : 0, own);
    if (trace) {
      std::chrono::steady_clock::time_point t2 = std::chrono::steady_clock::now();
      t_set += std::chrono::duration<double>(t1 - t0).count(); t_upd += std::chrono::duration<double>(t2 - t1).count();
      if (++n_calls % 256 == 0)
        fprintf(stderr, "[light_lof] %llu set_row: set_row+neighbours %.1f us, re-scoring of %zu rows %.1f us (means)\n",
                (unsigned long long)n_calls, t_set / n_calls * 1e6, order.size(), t_upd / n_calls * 1e6);
    }
    return true;
  }
  // LOF of a point from its neighbour list `i` of nb: lrd of the point over the mean lrd of the neighbours
  float score_from_neighbors(const raw_neighbors& nb, size_t i) const {
    const uint32_t cnt = nb.count.empty() ? 0 : nb.count[i];
    std::vector<float> lrds;
    float lrd;
    if (cnt == 0) {
      lrd = std::numeric_limits<float>::infinity();
    } else {
      lrds.resize(cnt);
      float sum_reach = 0.0f;
      for (uint32_t j = 0; j < cnt; ++j) {
        const uint64_t r = nb.idx[i * nb.k + j];
        if (r >= kdist_.size()) throw jubatus_exception("row not found in lof table: " + nn_->get_table()->get_key(r));
        lrds[j] = lrd_[r];
        sum_reach += std::max(nb.dist[i * nb.k + j], kdist_[r]);
      }
      lrd = sum_reach == 0 ? std::numeric_limits<float>::infinity() : cnt / sum_reach;
    }
    return calculate_lof(lrd, lrds);
  }
  std::shared_ptr<gpu_nearest_neighbor> nn_;
  int nn_num_, rnn_num_;
  std::shared_ptr<unlearner_base> unlearner_;
  std::vector<float> kdist_, lrd_;               // by row index of the engine's table
  std::unordered_set<uint64_t> param_dirty_;     // rows whose parameters changed here since the last MIX round
};

}  // namespace jubatus_b200
#endif  // JUBANN_HOST_HPP_
