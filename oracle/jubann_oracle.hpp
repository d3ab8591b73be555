// TEST INFRASTRUCTURE ONLY -- CPU oracle for the jubatus nearest_neighbor hot path.
//
// PARITY UNPINNED: this is a restatement of jubatus_core 0.3.2 (pinned by the
// reference at tools/packaging/rpm/package-config:12, linked at wscript:90), whose
// source is NOT under /root/reference and cannot be built or imported here.  The
// reference's own tests on this path (client_test/nearest_neighbor_test.cpp:57-98)
// assert no numeric result.  What *is* pinned: MT19937, FNV-1 and Jenkins mix64
// against their published known-answer vectors (tests/test_oracle_kat.py), and the
// rest by golden vectors generated from this file (tests/golden/).
// The uncertain details are frozen as listed in SURVEY.md A.9 and DESIGN.md.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may use this file.  The product (jubatus_b200/) never includes or links it.
//
// Each block cites the jubatus_core file it restates ([core] = jubatus_core 0.3.2,
// path relative to jubatus/core/) and the reference call site that reaches it.
#ifndef JUBANN_ORACLE_HPP_
#define JUBANN_ORACLE_HPP_

#include <stdint.h>
#include <math.h>
#include <float.h>
#include <string.h>

#include <algorithm>
#include <limits>
#include <map>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

namespace oracle {

typedef std::vector<std::pair<std::string, float> > sfv_t;  // [core] common/type.hpp

// ---------------------------------------------------------------------------
// [core] common/hash.hpp  hash_util::calc_string_hash  -- 64-bit FNV-1
// (multiply, then xor; `char` sign-extends).  SURVEY A.2.
// ---------------------------------------------------------------------------
inline uint64_t calc_string_hash(const char* s, size_t n) {
  uint64_t h = 14695981039346656037ULL;
  for (size_t i = 0; i < n; ++i) {
    h *= 1099511628211ULL;
    h ^= static_cast<uint64_t>(static_cast<int64_t>(static_cast<signed char>(s[i])));
  }
  return h;
}
inline uint64_t calc_string_hash(const std::string& s) {
  return calc_string_hash(s.data(), s.size());
}

// ---------------------------------------------------------------------------
// [core] jubatus/util/math/random.h + random/mersenne_twister.h (pficommon fork)
// MT19937: init_genrand / genrand_int32; next_double = res53;
// next_gaussian = Box-Muller, no cached second variate (4 draws each).
// Frozen form: sqrt(-2*log(1-a)) * cos(2*pi*b), a drawn first.
// ---------------------------------------------------------------------------
// ---------------------------------------------------------------------------
// Alternates for the recalled choices most likely to differ from jubatus_core, behind
// one switch each so that adopting the real behaviour is a flag, not a rewrite.  Defaults = the
// frozen choices.  The CUDA path has the same two switches (jubann_set_option "gaussian_form",
// "euclid_score"); tests/golden/nn_golden_variants.json freezes the oracle's answers under each.
//   gaussian_form  0: sqrt(-2 log(1-a)) * cos(2 pi b)   1: ... * sin(2 pi b)
//   euclid_score   0: cos in double, score expression in double, narrowed to float
//                  1: cosf(theta) and the whole score in float (std::cos on a float argument picks
//                     the float overload in C++)
// ---------------------------------------------------------------------------
struct variants {
  int gaussian_form;
  int euclid_score;
};
inline variants& variant_switches() {
  static variants v = {0, 0};
  return v;
}

class mtrand {
 public:
  explicit mtrand(uint32_t seed) { init(seed); }
  void init(uint32_t s) {
    mt_[0] = s;
    for (int i = 1; i < N; ++i) {
      mt_[i] = 1812433253U * (mt_[i - 1] ^ (mt_[i - 1] >> 30)) + static_cast<uint32_t>(i);
    }
    idx_ = N;
  }
  uint32_t next_int() {
    if (idx_ >= N) regen();
    uint32_t y = mt_[idx_++];
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680U;
    y ^= (y << 15) & 0xefc60000U;
    y ^= (y >> 18);
    return y;
  }
  double next_double() {
    uint32_t a = next_int() >> 5, b = next_int() >> 6;
    return (a * 67108864.0 + b) * (1.0 / 9007199254740992.0);
  }
  double next_gaussian() {
    double a = next_double();
    double b = next_double();
    if (variant_switches().gaussian_form == 1) return sqrt(-2.0 * log(1.0 - a)) * sin(2.0 * M_PI * b);
    return sqrt(-2.0 * log(1.0 - a)) * cos(2.0 * M_PI * b);
  }

 private:
  enum { N = 624, M = 397 };
  void regen() {
    static const uint32_t mag[2] = {0U, 0x9908b0dfU};
    int k = 0;
    for (; k < N - M; ++k) {
      uint32_t y = (mt_[k] & 0x80000000U) | (mt_[k + 1] & 0x7fffffffU);
      mt_[k] = mt_[k + M] ^ (y >> 1) ^ mag[y & 1U];
    }
    for (; k < N - 1; ++k) {
      uint32_t y = (mt_[k] & 0x80000000U) | (mt_[k + 1] & 0x7fffffffU);
      mt_[k] = mt_[k + (M - N)] ^ (y >> 1) ^ mag[y & 1U];
    }
    uint32_t y = (mt_[N - 1] & 0x80000000U) | (mt_[0] & 0x7fffffffU);
    mt_[N - 1] = mt_[M - 1] ^ (y >> 1) ^ mag[y & 1U];
    idx_ = 0;
  }
  uint32_t mt_[N];
  int idx_;
};

// ---------------------------------------------------------------------------
// [core] storage/bit_vector.hpp  bit_vector_base<uint64_t>   
// ---------------------------------------------------------------------------
inline size_t words_for_bits(size_t bits) { return (bits + 63) / 64; }

// SWAR bit count as bit_vector::bit_count does it (no hardware popcnt at -O2).
inline int bit_count(uint64_t bits) {
#if defined(__POPCNT__)
  return __builtin_popcountll(bits);
#else
  bits = (bits & 0x5555555555555555ULL) + ((bits >> 1) & 0x5555555555555555ULL);
  bits = (bits & 0x3333333333333333ULL) + ((bits >> 2) & 0x3333333333333333ULL);
  bits = (bits & 0x0f0f0f0f0f0f0f0fULL) + ((bits >> 4) & 0x0f0f0f0f0f0f0f0fULL);
  return static_cast<int>((bits * 0x0101010101010101ULL) >> 56);
#endif
}

struct bit_vector {
  size_t bit_num;
  std::vector<uint64_t> bits;
  explicit bit_vector(size_t n = 0) : bit_num(n), bits(words_for_bits(n), 0) {}
  void set_bit(size_t pos) {
    if (pos >= bit_num) throw std::out_of_range("set_bit");
    bits[pos / 64] |= (1ULL << (pos % 64));
  }
  bool get_bit(size_t pos) const { return (bits[pos / 64] >> (pos % 64)) & 1ULL; }
  uint64_t calc_hamming_distance(const bit_vector& o) const {
    if (bit_num != o.bit_num) throw std::invalid_argument("bit_num mismatch");
    uint64_t d = 0;
    for (size_t i = 0; i < bits.size(); ++i) d += bit_count(bits[i] ^ o.bits[i]);
    return d;
  }
};

inline uint32_t hamming_words(const uint64_t* a, const uint64_t* b, size_t w) {
  uint32_t d = 0;
  for (size_t i = 0; i < w; ++i) d += bit_count(a[i] ^ b[i]);
  return d;
}

// ---------------------------------------------------------------------------
// [core] nearest_neighbor/lsh_function.cpp  random_projection / binarize /
// cosine_lsh  (a4, a5).  Reached from
// jubatus/server/server/nearest_neighbor_serv.cpp:134 (set_row) and :153 (query).
// Frozen: proj[j] += v * (float)gaussian, float mul then float add, sfv order.
// ---------------------------------------------------------------------------
inline void projection_column(uint32_t seed, uint32_t hash_num, float* out) {
  mtrand rnd(seed);
  for (uint32_t j = 0; j < hash_num; ++j) out[j] = static_cast<float>(rnd.next_gaussian());
}

// Projection-vector cache keyed by the 32-bit seed (0.3.x cache).
class projection_cache {
 public:
  explicit projection_cache(uint32_t hash_num) : hash_num_(hash_num) {}
  const float* get(uint32_t seed) {
    std::unordered_map<uint32_t, std::vector<float> >::iterator it = map_.find(seed);
    if (it == map_.end()) {
      std::vector<float> col(hash_num_);
      projection_column(seed, hash_num_, col.data());
      it = map_.insert(std::make_pair(seed, col)).first;
    }
    return it->second.data();
  }
 private:
  uint32_t hash_num_;
  std::unordered_map<uint32_t, std::vector<float> > map_;
};

// random_projection over pre-hashed features (seed32[i], val[i]) in sfv order.
inline void random_projection(const uint32_t* seed32, const float* val, size_t nnz,
                              uint32_t hash_num, projection_cache* cache, float* proj) {
  for (uint32_t j = 0; j < hash_num; ++j) proj[j] = 0.0f;
  std::vector<float> tmp;
  for (size_t i = 0; i < nnz; ++i) {
    const float* g;
    if (cache) {
      g = cache->get(seed32[i]);
    } else {
      tmp.resize(hash_num);
      projection_column(seed32[i], hash_num, tmp.data());
      g = tmp.data();
    }
    const float v = val[i];
    for (uint32_t j = 0; j < hash_num; ++j) {
      volatile float prod = v * g[j];  // keep mul and add separately rounded
      proj[j] = proj[j] + prod;
    }
  }
}

inline void binarize(const float* proj, uint32_t hash_num, uint64_t* words) {
  const size_t w = words_for_bits(hash_num);
  for (size_t i = 0; i < w; ++i) words[i] = 0;
  for (uint32_t j = 0; j < hash_num; ++j) {
    if (proj[j] > 0.0f) words[j / 64] |= (1ULL << (j % 64));
  }
}

// [core] nearest_neighbor/euclid_lsh.cpp  calc_norm  (a6)
inline float calc_norm(const float* val, size_t nnz) {
  float sq = 0.0f;
  for (size_t i = 0; i < nnz; ++i) {
    volatile float p = val[i] * val[i];
    sq = sq + p;
  }
  return sqrtf(sq);
}

// ---------------------------------------------------------------------------
// [core] nearest_neighbor/minhash.cpp  hash_mix64 (Bob Jenkins lookup8.c mix64),
// calc_hash, minhash::hash  (a7)
// ---------------------------------------------------------------------------
inline void hash_mix64(uint64_t& a, uint64_t& b, uint64_t& c) {
  a -= b; a -= c; a ^= (c >> 43);
  b -= c; b -= a; b ^= (a << 9);
  c -= a; c -= b; c ^= (b >> 8);
  a -= b; a -= c; a ^= (c >> 38);
  b -= c; b -= a; b ^= (a << 23);
  c -= a; c -= b; c ^= (b >> 5);
  a -= b; a -= c; a ^= (c >> 35);
  b -= c; b -= a; b ^= (a << 49);
  c -= a; c -= b; c ^= (b >> 11);
  a -= b; a -= c; a ^= (c >> 12);
  b -= c; b -= a; b ^= (a << 18);
  c -= a; c -= b; c ^= (b >> 22);
}

inline float minhash_calc_hash(uint64_t a, uint64_t b, float val) {
  uint64_t c = 0xc3a5c85c97cb3127ULL;
  hash_mix64(a, b, c);
  hash_mix64(a, b, c);
  float r = static_cast<float>(a) / static_cast<float>(0xFFFFFFFFFFFFFFFFULL);
  return -logf(r) / val;
}

inline void minhash_signature(const uint64_t* keyhash64, const float* val, size_t nnz,
                              uint32_t hash_num, uint64_t* words, float* min_out = 0) {
  const size_t w = words_for_bits(hash_num);
  for (size_t i = 0; i < w; ++i) words[i] = 0;
  for (uint32_t j = 0; j < hash_num; ++j) {
    float best = FLT_MAX;
    uint64_t best_key = 0;
    for (size_t i = 0; i < nnz; ++i) {
      float hv = minhash_calc_hash(keyhash64[i], j, val[i]);
      if (hv < best) { best = hv; best_key = keyhash64[i]; }
    }
    if (min_out) min_out[j] = best;
    if (best_key & 1ULL) words[j / 64] |= (1ULL << (j % 64));
  }
}

// ---------------------------------------------------------------------------
// [core] storage/fixed_size_heap.hpp  (a11)
// bounded max-heap of the `limit` smallest values; get_sorted ascending.
// ---------------------------------------------------------------------------
template <typename T>
class fixed_size_heap {
 public:
  explicit fixed_size_heap(size_t limit) : limit_(limit) {}
  void push(const T& v) {
    if (limit_ == 0) return;
    if (data_.size() < limit_) {
      data_.push_back(v);
      std::push_heap(data_.begin(), data_.end());
    } else if (v < data_.front()) {
      std::pop_heap(data_.begin(), data_.end());
      data_.back() = v;
      std::push_heap(data_.begin(), data_.end());
    }
  }
  void get_sorted(std::vector<T>& out) const {
    out = data_;
    std::sort(out.begin(), out.end());
  }
  size_t size() const { return data_.size(); }
 private:
  size_t limit_;
  std::vector<T> data_;
};

// ---------------------------------------------------------------------------
// [core] nearest_neighbor/bit_vector_ranking.cpp  ranking_hamming_bit_vectors
// (a10).  Rows are `w` u64 words each, contiguous.  `threads`
// partitions rows in contiguous chunks and merges per-thread heaps; the result is
// identical to the single-thread result because (dist, index) is a total order.
// Output: ascending (dist, row index); score = dist / (float)hash_num.
// ---------------------------------------------------------------------------
typedef std::pair<uint32_t, uint64_t> dist_idx;

inline void ranking_hamming_range(const uint64_t* q, const uint64_t* rows, size_t w,
                                  uint64_t begin, uint64_t end, fixed_size_heap<dist_idx>& heap) {
  for (uint64_t i = begin; i < end; ++i) {
    heap.push(dist_idx(hamming_words(q, rows + i * w, w), i));
  }
}

inline void ranking_hamming_bit_vectors(const uint64_t* q, const uint64_t* rows, uint64_t n,
                                        uint32_t hash_num, uint64_t ret_num, unsigned threads,
                                        std::vector<std::pair<uint64_t, float> >& ret) {
  const size_t w = words_for_bits(hash_num);
  std::vector<dist_idx> sorted;
  if (threads <= 1 || n < 2 * static_cast<uint64_t>(threads)) {
    fixed_size_heap<dist_idx> heap(ret_num);
    ranking_hamming_range(q, rows, w, 0, n, heap);
    heap.get_sorted(sorted);
  } else {
    std::vector<fixed_size_heap<dist_idx> > heaps(threads, fixed_size_heap<dist_idx>(ret_num));
    std::vector<std::thread> th;
    const uint64_t chunk = (n + threads - 1) / threads;
    for (unsigned t = 0; t < threads; ++t) {
      uint64_t b = std::min<uint64_t>(n, t * chunk), e = std::min<uint64_t>(n, b + chunk);
      th.push_back(std::thread(ranking_hamming_range, q, rows, w, b, e, std::ref(heaps[t])));
    }
    for (size_t t = 0; t < th.size(); ++t) th[t].join();
    fixed_size_heap<dist_idx> heap(ret_num);
    for (unsigned t = 0; t < threads; ++t) {
      std::vector<dist_idx> part;
      heaps[t].get_sorted(part);
      for (size_t i = 0; i < part.size(); ++i) heap.push(part[i]);
    }
    heap.get_sorted(sorted);
  }
  ret.clear();
  const float denom = static_cast<float>(hash_num);
  for (size_t i = 0; i < sorted.size(); ++i) {
    ret.push_back(std::make_pair(sorted[i].second, static_cast<float>(sorted[i].first) / denom));
  }
}

// ---------------------------------------------------------------------------
// [core] nearest_neighbor/euclid_lsh.cpp  neighbor_row_from_hash  (a13)
// theta is held in float, cos evaluated in double (frozen, A.9); the running
// expression is double (float operands promoted), narrowed to float for the heap.
// Heap orders (score, index); output sqrt(norm_q^2 + score), no clamp.
// ---------------------------------------------------------------------------
inline float euclid_score(float norm_i, float norm_q, uint32_t hamm, uint32_t hash_num) {
  const float denom = static_cast<float>(hash_num);
  const float theta = static_cast<float>(hamm * M_PI / denom);
  if (variant_switches().euclid_score == 1) {   // every operation in float, one rounding each
    volatile float cf = cosf(theta);
    volatile float tf = (2 * norm_q) * cf;
    volatile float uf = norm_i - tf;
    volatile float sf = norm_i * uf;
    return sf;
  }
  volatile double c = cos(static_cast<double>(theta));
  volatile double t = static_cast<double>(2 * norm_q) * c;
  volatile double u = static_cast<double>(norm_i) - t;
  volatile double s = static_cast<double>(norm_i) * u;
  return static_cast<float>(s);
}

typedef std::pair<float, uint64_t> score_idx;

inline void euclid_range(const uint64_t* q, float norm_q, const uint64_t* rows, const float* norms,
                         size_t w, uint32_t hash_num, uint64_t begin, uint64_t end,
                         fixed_size_heap<score_idx>& heap) {
  for (uint64_t i = begin; i < end; ++i) {
    uint32_t h = hamming_words(q, rows + i * w, w);
    heap.push(score_idx(euclid_score(norms[i], norm_q, h, hash_num), i));
  }
}

inline void ranking_euclid(const uint64_t* q, float norm_q, const uint64_t* rows, const float* norms,
                           uint64_t n, uint32_t hash_num, uint64_t ret_num, unsigned threads,
                           std::vector<std::pair<uint64_t, float> >& ret) {
  const size_t w = words_for_bits(hash_num);
  std::vector<score_idx> sorted;
  if (threads <= 1 || n < 2 * static_cast<uint64_t>(threads)) {
    fixed_size_heap<score_idx> heap(ret_num);
    euclid_range(q, norm_q, rows, norms, w, hash_num, 0, n, heap);
    heap.get_sorted(sorted);
  } else {
    std::vector<fixed_size_heap<score_idx> > heaps(threads, fixed_size_heap<score_idx>(ret_num));
    std::vector<std::thread> th;
    const uint64_t chunk = (n + threads - 1) / threads;
    for (unsigned t = 0; t < threads; ++t) {
      uint64_t b = std::min<uint64_t>(n, t * chunk), e = std::min<uint64_t>(n, b + chunk);
      th.push_back(std::thread(euclid_range, q, norm_q, rows, norms, w, hash_num, b, e,
                               std::ref(heaps[t])));
    }
    for (size_t t = 0; t < th.size(); ++t) th[t].join();
    fixed_size_heap<score_idx> heap(ret_num);
    for (unsigned t = 0; t < threads; ++t) {
      std::vector<score_idx> part;
      heaps[t].get_sorted(part);
      for (size_t i = 0; i < part.size(); ++i) heap.push(part[i]);
    }
    heap.get_sorted(sorted);
  }
  ret.clear();
  for (size_t i = 0; i < sorted.size(); ++i) {
    volatile float nq2 = norm_q * norm_q;
    volatile float s = nq2 + sorted[i].first;
    ret.push_back(std::make_pair(sorted[i].second, sqrtf(s)));
  }
}

// ---------------------------------------------------------------------------
// [core] storage/column_table.hpp  (a9) -- the parts the NN path uses:
// one bit_vector column [+ one float column], keys, versions (owner, clock),
// key->index map, add (append or in-place update), delete_row (swap with last),
// exact_match, get_key, set_row-from-diff ("newer or equal clock wins").
// ---------------------------------------------------------------------------
typedef std::pair<std::string, uint64_t> version_t;  // (owner, clock)

struct packed_row {  // what column_table::get_row serialises (msgpack in the reference)
  std::string key;
  version_t version;
  std::vector<uint64_t> sig;
  std::vector<float> floats;  // euclid_lsh: [norm]; light_lof table: [kdist, lrd]
};

class column_table {
 public:
  column_table() : hash_num_(0), nfloat_(0), clock_(0) {}
  void init(uint32_t hash_num, size_t nfloat) {
    hash_num_ = hash_num; nfloat_ = nfloat; w_ = words_for_bits(hash_num);
    clear();
  }
  void clear() {
    keys_.clear(); versions_.clear(); index_.clear(); sigs_.clear(); floats_.clear(); clock_ = 0;
  }
  uint64_t size() const { return keys_.size(); }
  uint32_t hash_num() const { return hash_num_; }
  size_t words() const { return w_; }
  size_t nfloat() const { return nfloat_; }
  // returns true iff the key is new
  bool add(const std::string& key, const std::string& owner, const uint64_t* sig, const float* fl) {
    std::unordered_map<std::string, uint64_t>::iterator it = index_.find(key);
    bool is_new = (it == index_.end());
    uint64_t idx;
    if (is_new) {
      idx = keys_.size();
      keys_.push_back(key);
      versions_.push_back(version_t(owner, clock_));
      index_[key] = idx;
      sigs_.resize(sigs_.size() + w_);
      floats_.resize(floats_.size() + nfloat_);
    } else {
      idx = it->second;
      versions_[idx] = version_t(owner, clock_);
    }
    for (size_t i = 0; i < w_; ++i) sigs_[idx * w_ + i] = sig ? sig[i] : 0;
    for (size_t i = 0; i < nfloat_; ++i) floats_[idx * nfloat_ + i] = fl ? fl[i] : 0.0f;
    ++clock_;
    return is_new;
  }
  bool delete_row(const std::string& key) {
    std::unordered_map<std::string, uint64_t>::iterator it = index_.find(key);
    if (it == index_.end()) return false;
    const uint64_t idx = it->second, last = keys_.size() - 1;
    if (idx != last) {
      keys_[idx] = keys_[last];
      versions_[idx] = versions_[last];
      for (size_t i = 0; i < w_; ++i) sigs_[idx * w_ + i] = sigs_[last * w_ + i];
      for (size_t i = 0; i < nfloat_; ++i) floats_[idx * nfloat_ + i] = floats_[last * nfloat_ + i];
      index_[keys_[idx]] = idx;
    }
    index_.erase(it);
    keys_.pop_back(); versions_.pop_back();
    sigs_.resize(last * w_); floats_.resize(last * nfloat_);
    return true;
  }
  std::pair<bool, uint64_t> exact_match(const std::string& key) const {
    std::unordered_map<std::string, uint64_t>::const_iterator it = index_.find(key);
    if (it == index_.end()) return std::make_pair(false, 0);
    return std::make_pair(true, it->second);
  }
  const std::string& get_key(uint64_t idx) const { return keys_.at(idx); }
  const version_t& get_version(uint64_t idx) const { return versions_.at(idx); }
  const uint64_t* sig(uint64_t idx) const { return &sigs_[idx * w_]; }
  const uint64_t* sigs() const { return sigs_.data(); }
  float get_float(uint64_t idx, size_t c) const { return floats_[idx * nfloat_ + c]; }
  void set_float(uint64_t idx, size_t c, float v) { floats_[idx * nfloat_ + c] = v; }
  const float* floats() const { return floats_.data(); }
  uint64_t clock() const { return clock_; }
  packed_row get_row(uint64_t idx) const {
    packed_row r;
    r.key = keys_.at(idx); r.version = versions_[idx];
    r.sig.assign(sig(idx), sig(idx) + w_);
    r.floats.assign(&floats_[idx * nfloat_], &floats_[idx * nfloat_] + nfloat_);
    return r;
  }
  // column_table::set_row(packed): upsert by key; existing row is overwritten only if
  // its clock <= incoming clock; local clock moves past the incoming one.
  version_t set_row(const packed_row& r) {
    std::unordered_map<std::string, uint64_t>::iterator it = index_.find(r.key);
    if (it == index_.end()) {
      uint64_t idx = keys_.size();
      keys_.push_back(r.key); versions_.push_back(r.version); index_[r.key] = idx;
      sigs_.insert(sigs_.end(), r.sig.begin(), r.sig.end());
      floats_.insert(floats_.end(), r.floats.begin(), r.floats.end());
    } else {
      uint64_t idx = it->second;
      if (versions_[idx].second <= r.version.second) {
        versions_[idx] = r.version;
        for (size_t i = 0; i < w_; ++i) sigs_[idx * w_ + i] = r.sig[i];
        for (size_t i = 0; i < nfloat_; ++i) floats_[idx * nfloat_ + i] = r.floats[i];
      }
    }
    if (clock_ <= r.version.second) clock_ = r.version.second + 1;
    return r.version;
  }
  std::vector<std::string> get_all_row_ids() const { return keys_; }

 private:
  uint32_t hash_num_;
  size_t w_, nfloat_;
  uint64_t clock_;
  std::vector<std::string> keys_;
  std::vector<version_t> versions_;
  std::unordered_map<std::string, uint64_t> index_;
  std::vector<uint64_t> sigs_;
  std::vector<float> floats_;
};

// ---------------------------------------------------------------------------
// [core] framework/mixable_versioned_table.{hpp,cpp}  (a17)
// driven by jubatus/server/framework/mixer/push_mixer.cpp:410-472 (pull /
// get_pull_argument / push) and linear_mixer.cpp:547-561,613-628 (get_diff/put_diff).
// Frozen: get_argument folds the local rows' versions into the clock first.
// ---------------------------------------------------------------------------
typedef std::map<std::string, uint64_t> version_clock;  // owner -> max clock seen

class mixable_versioned_table {
 public:
  explicit mixable_versioned_table(column_table* t) : table_(t) {}
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
    diff.clear();
    for (uint64_t i = 0; i < table_->size(); ++i) {
      const version_t& v = table_->get_version(i);
      version_clock::const_iterator it = arg.find(v.first);
      if (it == arg.end() || it->second < v.second) diff.push_back(table_->get_row(i));
    }
  }
  void push(const std::vector<packed_row>& diff) {
    for (size_t i = 0; i < diff.size(); ++i) update_version(table_->set_row(diff[i]));
  }
  // linear_mixable shape: get_diff = everything newer than the last put; mix = concat;
  // put_diff = push.
  void get_diff(std::vector<packed_row>& diff) const { pull(vc_, diff); }
  static void mix(const std::vector<packed_row>& lhs, std::vector<packed_row>& acc) {
    acc.insert(acc.end(), lhs.begin(), lhs.end());
  }
  bool put_diff(const std::vector<packed_row>& diff) { push(diff); return true; }

 private:
  column_table* table_;
  version_clock vc_;
};

// ---------------------------------------------------------------------------
// [core] nearest_neighbor/{nearest_neighbor_base,bit_vector_nearest_neighbor_base,
// lsh,euclid_lsh,minhash}.cpp + nearest_neighbor_factory.cpp  (a12-a14)
// Constructed at jubatus/server/server/nearest_neighbor_serv.cpp:106-108.
// ---------------------------------------------------------------------------
enum method_t { METHOD_LSH = 0, METHOD_EUCLID_LSH = 1, METHOD_MINHASH = 2 };

inline method_t parse_method(const std::string& m) {
  if (m == "lsh") return METHOD_LSH;
  if (m == "euclid_lsh") return METHOD_EUCLID_LSH;
  if (m == "minhash") return METHOD_MINHASH;
  throw std::invalid_argument("unsupported_method: " + m);
}

// pre-hashed sparse feature vector: what the engine derives from sfv_t keys
struct hashed_sfv {
  std::vector<uint32_t> seed32;
  std::vector<uint64_t> keyhash64;
  std::vector<float> val;
};
inline hashed_sfv hash_sfv(const sfv_t& sfv) {
  hashed_sfv h;
  for (size_t i = 0; i < sfv.size(); ++i) {
    uint64_t k = calc_string_hash(sfv[i].first);
    h.keyhash64.push_back(k);
    h.seed32.push_back(static_cast<uint32_t>(k));
    h.val.push_back(sfv[i].second);
  }
  return h;
}

class nearest_neighbor {
 public:
  nearest_neighbor(const std::string& method, int hash_num, const std::string& my_id,
                   unsigned threads = 1)
      : method_(parse_method(method)), hash_num_(hash_num), my_id_(my_id), threads_(threads),
        cache_(hash_num < 1 ? 1 : hash_num), mixable_(&table_) {
    if (hash_num < 1) throw std::invalid_argument("invalid_parameter: hash_num should be >= 1");
    table_.init(hash_num, method_ == METHOD_EUCLID_LSH ? 1 : 0);
  }
  std::string type() const {
    return method_ == METHOD_LSH ? "lsh" : method_ == METHOD_EUCLID_LSH ? "euclid_lsh" : "minhash";
  }
  void clear() { table_.clear(); }
  uint32_t hash_num() const { return hash_num_; }
  method_t method() const { return method_; }
  column_table* get_table() { return &table_; }
  const column_table* get_const_table() const { return &table_; }
  mixable_versioned_table* get_mixable() { return &mixable_; }

  // signature (+ norm for euclid_lsh) of a pre-hashed sparse vector
  void hash(const hashed_sfv& h, std::vector<uint64_t>& sig, float& norm) {
    sig.assign(table_.words(), 0);
    norm = 0.0f;
    if (method_ == METHOD_MINHASH) {
      minhash_signature(h.keyhash64.data(), h.val.data(), h.val.size(), hash_num_, sig.data());
    } else {
      std::vector<float> proj(hash_num_);
      random_projection(h.seed32.data(), h.val.data(), h.val.size(), hash_num_, &cache_, proj.data());
      binarize(proj.data(), hash_num_, sig.data());
      if (method_ == METHOD_EUCLID_LSH) norm = calc_norm(h.val.data(), h.val.size());
    }
  }
  void set_row(const std::string& id, const sfv_t& sfv) {
    std::vector<uint64_t> sig; float norm;
    hash(hash_sfv(sfv), sig, norm);
    table_.add(id, my_id_, sig.data(), &norm);
  }
  void neighbor_row_from_hash(const uint64_t* sig, float norm, uint64_t ret_num,
                              std::vector<std::pair<std::string, float> >& out) const {
    std::vector<std::pair<uint64_t, float> > r;
    if (method_ == METHOD_EUCLID_LSH) {
      ranking_euclid(sig, norm, table_.sigs(), table_.floats(), table_.size(), hash_num_, ret_num,
                     threads_, r);
    } else {
      ranking_hamming_bit_vectors(sig, table_.sigs(), table_.size(), hash_num_, ret_num, threads_, r);
    }
    out.clear();
    for (size_t i = 0; i < r.size(); ++i) out.push_back(std::make_pair(table_.get_key(r[i].first), r[i].second));
  }
  void neighbor_row(const sfv_t& q, std::vector<std::pair<std::string, float> >& out, uint64_t ret_num) {
    std::vector<uint64_t> sig; float norm;
    hash(hash_sfv(q), sig, norm);
    neighbor_row_from_hash(sig.data(), norm, ret_num, out);
  }
  void neighbor_row(const std::string& id, std::vector<std::pair<std::string, float> >& out,
                    uint64_t ret_num) const {
    std::pair<bool, uint64_t> hit = table_.exact_match(id);
    if (!hit.first) { out.clear(); return; }
    float norm = table_.nfloat() ? table_.get_float(hit.second, 0) : 0.0f;
    neighbor_row_from_hash(table_.sig(hit.second), norm, ret_num, out);
  }
  float calc_similarity(float d) const { return method_ == METHOD_EUCLID_LSH ? -d : 1.0f - d; }
  template <typename Q>
  void similar_row(const Q& q, std::vector<std::pair<std::string, float> >& out, uint64_t ret_num) {
    neighbor_row(q, out, ret_num);
    for (size_t i = 0; i < out.size(); ++i) out[i].second = calc_similarity(out[i].second);
  }
  std::vector<std::string> get_all_row_ids() const { return table_.get_all_row_ids(); }

 private:
  method_t method_;
  uint32_t hash_num_;
  std::string my_id_;
  unsigned threads_;
  column_table table_;
  projection_cache cache_;
  mixable_versioned_table mixable_;
};

// ---------------------------------------------------------------------------
// [core] anomaly/light_lof.cpp  (a16) reached from
// jubatus/server/server/anomaly_serv.cpp:134-138 (factory), :169 (add), :238 (calc_score).
// Second table: columns (kdist, lrd) keyed by the same ids.
// ---------------------------------------------------------------------------
class light_lof {
 public:
  light_lof(nearest_neighbor* nn, int nn_num, int rnn_num, const std::string& my_id)
      : nn_(nn), nn_num_(nn_num), rnn_num_(rnn_num), my_id_(my_id) {
    if (nn_num < 2) throw std::invalid_argument("nearest_neighbor_num should be >= 2");
    if (rnn_num < nn_num) throw std::invalid_argument("reverse_nearest_neighbor_num should be >= nearest_neighbor_num");
    lof_.init(1, 2);  // signature column unused; floats = (kdist, lrd)
  }
  void clear() { nn_->clear(); lof_.clear(); }
  static float calculate_lof(float lrd, const std::vector<float>& neighbor_lrds) {
    if (neighbor_lrds.empty()) return lrd == 0 ? 1.0f : std::numeric_limits<float>::infinity();
    float sum = 0.0f;
    for (size_t i = 0; i < neighbor_lrds.size(); ++i) sum += neighbor_lrds[i];
    if (std::isinf(sum) && std::isinf(lrd)) return 1.0f;
    return sum / (neighbor_lrds.size() * lrd);
  }
  float collect_lrds_from_neighbors(const std::vector<std::pair<std::string, float> >& nb,
                                    std::vector<float>& lrds) const {
    if (nb.empty()) return std::numeric_limits<float>::infinity();
    lrds.resize(nb.size());
    std::vector<float> kd(nb.size());
    for (size_t i = 0; i < nb.size(); ++i) {
      std::pair<bool, uint64_t> hit = lof_.exact_match(nb[i].first);
      if (!hit.first) throw std::runtime_error("row not found in lof table: " + nb[i].first);
      kd[i] = lof_.get_float(hit.second, 0);
      lrds[i] = lof_.get_float(hit.second, 1);
    }
    float sum_reach = 0.0f;
    for (size_t i = 0; i < nb.size(); ++i) sum_reach += std::max(nb[i].second, kd[i]);
    if (sum_reach == 0) return std::numeric_limits<float>::infinity();
    return nb.size() / sum_reach;
  }
  float calc_anomaly_score(const sfv_t& q) {
    std::vector<std::pair<std::string, float> > nb;
    nn_->neighbor_row(q, nb, nn_num_);
    std::vector<float> lrds;
    float lrd = collect_lrds_from_neighbors(nb, lrds);
    return calculate_lof(lrd, lrds);
  }
  float calc_anomaly_score(const std::string& id) {
    std::vector<std::pair<std::string, float> > nb;
    nn_->neighbor_row(id, nb, nn_num_);
    std::vector<float> lrds;
    float lrd = collect_lrds_from_neighbors(nb, lrds);
    return calculate_lof(lrd, lrds);
  }
  void collect_neighbors(const std::string& id, std::vector<std::string>& order,
                         std::unordered_set<std::string>& set) const {
    std::vector<std::pair<std::string, float> > nb;
    nn_->neighbor_row(id, nb, rnn_num_);
    for (size_t i = 0; i < nb.size(); ++i)
      if (set.insert(nb[i].first).second) order.push_back(nb[i].first);
  }
  bool set_row(const std::string& id, const sfv_t& sfv) {
    std::vector<std::string> order;
    std::unordered_set<std::string> set;
    if (nn_->get_table()->exact_match(id).first) collect_neighbors(id, order, set);
    nn_->set_row(id, sfv);
    collect_neighbors(id, order, set);
    float dummy[2] = {-1.0f, -1.0f};
    lof_.add(id, my_id_, 0, dummy);
    if (set.insert(id).second) order.push_back(id);
    update_entries(order);
    return true;
  }
  void update_entries(const std::vector<std::string>& ids) {
    std::vector<std::vector<std::pair<std::string, float> > > nn(ids.size());
    std::vector<uint64_t> row(ids.size());
    for (size_t i = 0; i < ids.size(); ++i) {
      nn_->neighbor_row(ids[i], nn[i], nn_num_);
      std::pair<bool, uint64_t> hit = lof_.exact_match(ids[i]);
      if (!hit.first) throw std::runtime_error("row not found in lof table: " + ids[i]);
      row[i] = hit.second;
      lof_.set_float(row[i], 0, nn[i].empty() ? 0.0f : nn[i].back().second);
    }
    for (size_t i = 0; i < ids.size(); ++i) {
      const size_t length = std::min<size_t>(nn[i].size(), nn_num_);
      float sum_reach = 0.0f;
      for (size_t j = 0; j < length; ++j) {
        std::pair<bool, uint64_t> hit = lof_.exact_match(nn[i][j].first);
        float kd = hit.first ? lof_.get_float(hit.second, 0) : 0.0f;
        sum_reach += std::max(nn[i][j].second, kd);
      }
      if (sum_reach == 0) { lof_.set_float(row[i], 1, std::numeric_limits<float>::infinity()); continue; }
      lof_.set_float(row[i], 1, length / sum_reach);
    }
  }
  column_table* lof_table() { return &lof_; }
  nearest_neighbor* engine() { return nn_; }

 private:
  nearest_neighbor* nn_;
  int nn_num_, rnn_num_;
  std::string my_id_;
  column_table lof_;
};

// ---------------------------------------------------------------------------
// Synthetic workload of SURVEY.md 8(d): splitmix64, master seed 0x4A554241.
// (Not a restatement of the reference: shared input generator for CPU and GPU.)
// ---------------------------------------------------------------------------
inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}
inline uint64_t synth_draw(uint64_t seed, uint64_t row, uint64_t t) {
  return splitmix64(splitmix64(seed ^ splitmix64(row)) + t);
}
// Row `row`: first `nnz` distinct ids among draws t=0,1,.. (id = draw & (nfeat-1)), sorted
// ascending; value of id = 1 + ((draw(row, 2^32 + id) >> 40) / 2^24).
inline void synth_row(uint64_t seed, uint64_t row, uint32_t nnz, uint32_t nfeat_log2,
                      uint32_t* ids, float* vals) {
  const uint32_t mask = (1U << nfeat_log2) - 1U;
  uint32_t got = 0;
  for (uint64_t t = 0; got < nnz; ++t) {
    uint32_t id = static_cast<uint32_t>(synth_draw(seed, row, t)) & mask;
    bool dup = false;
    for (uint32_t i = 0; i < got; ++i) if (ids[i] == id) { dup = true; break; }
    if (!dup) ids[got++] = id;
  }
  std::sort(ids, ids + nnz);
  for (uint32_t i = 0; i < nnz; ++i) {
    uint64_t d = synth_draw(seed, row, (1ULL << 32) + ids[i]);
    vals[i] = 1.0f + static_cast<float>(d >> 40) * (1.0f / 16777216.0f);
  }
}
// feature id f -> key string "<f>@num" (what the num rule emits for datum key "<f>")
inline std::string synth_key(uint32_t f) { return std::to_string(f) + "@num"; }

}  // namespace oracle
#endif  // JUBANN_ORACLE_HPP_
