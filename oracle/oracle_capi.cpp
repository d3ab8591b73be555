// TEST INFRASTRUCTURE ONLY -- flat C entry points over oracle/jubann_oracle.hpp so that
// tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs can drive the oracle
// through ctypes.  PARITY UNPINNED (see the header of jubann_oracle.hpp).
// Never linked into, loaded by, or called from the product library.
#include "jubann_oracle.hpp"

#include <chrono>
#include <cstdio>
#include <memory>

using namespace oracle;

namespace {
struct engine_box {
  std::unique_ptr<nearest_neighbor> nn;
  std::unique_ptr<light_lof> lof;
  std::string err;
};
thread_local std::string g_err;

sfv_t make_sfv(const char* const* keys, const float* vals, uint64_t n) {
  sfv_t s;
  for (uint64_t i = 0; i < n; ++i) s.push_back(std::make_pair(std::string(keys[i]), vals[i]));
  return s;
}
int emit(const std::vector<std::pair<std::string, float> >& r, char* ids_buf, uint64_t ids_cap,
         float* scores, uint64_t* count) {
  // ids are written '\n'-separated into ids_buf
  std::string joined;
  for (size_t i = 0; i < r.size(); ++i) { joined += r[i].first; joined += '\n'; scores[i] = r[i].second; }
  if (joined.size() + 1 > ids_cap) { g_err = "ids buffer too small"; return -1; }
  memcpy(ids_buf, joined.c_str(), joined.size() + 1);
  *count = r.size();
  return 0;
}
}  // namespace

extern "C" {

const char* oracle_last_error() { return g_err.c_str(); }
// the alternates of SURVEY A.9 (see `variants` in jubann_oracle.hpp); -1 leaves a switch as it is
void oracle_set_variants(int gaussian_form, int euclid_score) {
  if (gaussian_form >= 0) variant_switches().gaussian_form = gaussian_form;
  if (euclid_score >= 0) variant_switches().euclid_score = euclid_score;
}

uint64_t oracle_fnv1_64(const char* s, uint64_t n) { return calc_string_hash(s, n); }

void oracle_mt_ints(uint32_t seed, uint64_t n, uint32_t* out) {
  mtrand r(seed);
  for (uint64_t i = 0; i < n; ++i) out[i] = r.next_int();
}
void oracle_mt_doubles(uint32_t seed, uint64_t n, double* out) {
  mtrand r(seed);
  for (uint64_t i = 0; i < n; ++i) out[i] = r.next_double();
}
void oracle_mt_gaussians(uint32_t seed, uint64_t n, double* out) {
  mtrand r(seed);
  for (uint64_t i = 0; i < n; ++i) out[i] = r.next_gaussian();
}
void oracle_mix64(uint64_t* abc) { hash_mix64(abc[0], abc[1], abc[2]); }
float oracle_minhash_calc_hash(uint64_t a, uint64_t b, float v) { return minhash_calc_hash(a, b, v); }

void oracle_projection_column(uint32_t seed, uint32_t hash_num, float* out) {
  projection_column(seed, hash_num, out);
}

// Signatures for a CSR batch of pre-hashed sparse vectors.  method: 0 lsh, 1 euclid_lsh,
// 2 minhash.  out_aux (optional): n*hash_num floats = projections (lsh/euclid) or the
// winning min-hash values (minhash), so tests can bound disagreements by |projection|.
int oracle_signatures_csr(int method, uint32_t hash_num, uint64_t n, const uint64_t* indptr,
                          const uint32_t* seed32, const uint64_t* keyhash64, const float* vals,
                          uint64_t* out_sigs, float* out_norms, float* out_aux, unsigned threads) {
  if (hash_num < 1) { g_err = "hash_num should be >= 1"; return -1; }
  const size_t w = words_for_bits(hash_num);
  if (threads < 1) threads = 1;
  std::vector<std::thread> th;
  for (unsigned t = 0; t < threads; ++t) {
    th.push_back(std::thread([=]() {
      projection_cache cache(hash_num);
      std::vector<float> proj(hash_num);
      for (uint64_t r = t; r < n; r += threads) {
        const uint64_t b = indptr[r], e = indptr[r + 1];
        if (method == METHOD_MINHASH) {
          minhash_signature(keyhash64 + b, vals + b, e - b, hash_num, out_sigs + r * w,
                            out_aux ? out_aux + r * hash_num : 0);
        } else {
          random_projection(seed32 + b, vals + b, e - b, hash_num, &cache, proj.data());
          binarize(proj.data(), hash_num, out_sigs + r * w);
          if (out_aux) memcpy(out_aux + r * hash_num, proj.data(), hash_num * sizeof(float));
        }
        if (out_norms) out_norms[r] = (method == METHOD_EUCLID_LSH) ? calc_norm(vals + b, e - b) : 0.0f;
      }
    }));
  }
  for (size_t t = 0; t < th.size(); ++t) th[t].join();
  return 0;
}

// One query against a contiguous row-major table.  Returns count written.
uint64_t oracle_ranking_hamming(const uint64_t* q, const uint64_t* rows, uint64_t n, uint32_t hash_num,
                                uint64_t ret_num, unsigned threads, uint64_t* out_idx, float* out_score) {
  std::vector<std::pair<uint64_t, float> > r;
  ranking_hamming_bit_vectors(q, rows, n, hash_num, ret_num, threads, r);
  for (size_t i = 0; i < r.size(); ++i) { out_idx[i] = r[i].first; out_score[i] = r[i].second; }
  return r.size();
}
uint64_t oracle_ranking_euclid(const uint64_t* q, float norm_q, const uint64_t* rows, const float* norms,
                               uint64_t n, uint32_t hash_num, uint64_t ret_num, unsigned threads,
                               uint64_t* out_idx, float* out_score) {
  std::vector<std::pair<uint64_t, float> > r;
  ranking_euclid(q, norm_q, rows, norms, n, hash_num, ret_num, threads, r);
  for (size_t i = 0; i < r.size(); ++i) { out_idx[i] = r[i].first; out_score[i] = r[i].second; }
  return r.size();
}

// Batch of queries, one after another (each using `threads` row partitions, the
// reference's own multi-thread mode).  out_* are [nq][ret_num]; out_count [nq].
// similar != 0 applies calc_similarity.  Returns elapsed seconds of the scan loop.
double oracle_scan_batch(int method, uint64_t nq, const uint64_t* qsigs, const float* qnorms,
                         const uint64_t* rows, const float* norms, uint64_t n, uint32_t hash_num,
                         uint64_t ret_num, unsigned threads, int similar,
                         uint64_t* out_idx, float* out_score, uint64_t* out_count) {
  const size_t w = words_for_bits(hash_num);
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  for (uint64_t qi = 0; qi < nq; ++qi) {
    std::vector<std::pair<uint64_t, float> > r;
    if (method == METHOD_EUCLID_LSH)
      ranking_euclid(qsigs + qi * w, qnorms[qi], rows, norms, n, hash_num, ret_num, threads, r);
    else
      ranking_hamming_bit_vectors(qsigs + qi * w, rows, n, hash_num, ret_num, threads, r);
    for (size_t i = 0; i < r.size(); ++i) {
      out_idx[qi * ret_num + i] = r[i].first;
      float s = r[i].second;
      if (similar) s = (method == METHOD_EUCLID_LSH) ? -s : 1.0f - s;
      out_score[qi * ret_num + i] = s;
    }
    out_count[qi] = r.size();
  }
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// Synthetic rows [row0, row0+n): ids/vals are [n][nnz].
void oracle_synth_rows(uint64_t seed, uint64_t row0, uint64_t n, uint32_t nnz, uint32_t nfeat_log2,
                       uint32_t* ids, float* vals, unsigned threads) {
  if (threads < 1) threads = 1;
  std::vector<std::thread> th;
  for (unsigned t = 0; t < threads; ++t)
    th.push_back(std::thread([=]() {
      for (uint64_t r = t; r < n; r += threads)
        synth_row(seed, row0 + r, nnz, nfeat_log2, ids + r * nnz, vals + r * nnz);
    }));
  for (size_t t = 0; t < th.size(); ++t) th[t].join();
}
// seed32 / keyhash64 tables for feature ids [0, nfeat): key "<f>@num"
void oracle_synth_feature_tables(uint32_t nfeat, uint32_t* seed32, uint64_t* keyhash64) {
  for (uint32_t f = 0; f < nfeat; ++f) {
    uint64_t k = calc_string_hash(synth_key(f));
    if (keyhash64) keyhash64[f] = k;
    if (seed32) seed32[f] = static_cast<uint32_t>(k);
  }
}
// Uniform random signature words (scan-only microbenchmarks; labelled as such).
void oracle_fill_random_words(uint64_t seed, uint64_t nwords, uint64_t* out, unsigned threads) {
  if (threads < 1) threads = 1;
  std::vector<std::thread> th;
  for (unsigned t = 0; t < threads; ++t)
    th.push_back(std::thread([=]() {
      for (uint64_t i = t; i < nwords; i += threads) out[i] = splitmix64(seed + i);
    }));
  for (size_t t = 0; t < th.size(); ++t) th[t].join();
}

// ---- engine-level surface (string ids), mirroring the driver/RPC names ----
void* oracle_nn_create(const char* method, int hash_num, const char* my_id, unsigned threads) {
  try {
    engine_box* b = new engine_box;
    b->nn.reset(new nearest_neighbor(method, hash_num, my_id, threads));
    return b;
  } catch (const std::exception& e) { g_err = e.what(); return 0; }
}
void* oracle_lof_create(const char* method, int hash_num, int nn_num, int rnn_num, const char* my_id) {
  try {
    engine_box* b = new engine_box;
    b->nn.reset(new nearest_neighbor(method, hash_num, my_id, 1));
    b->lof.reset(new light_lof(b->nn.get(), nn_num, rnn_num, my_id));
    return b;
  } catch (const std::exception& e) { g_err = e.what(); return 0; }
}
void oracle_nn_destroy(void* h) { delete static_cast<engine_box*>(h); }
void oracle_nn_clear(void* h) {
  engine_box* b = static_cast<engine_box*>(h);
  if (b->lof) b->lof->clear(); else b->nn->clear();
}
uint64_t oracle_nn_size(void* h) { return static_cast<engine_box*>(h)->nn->get_table()->size(); }
int oracle_nn_set_row(void* h, const char* id, const char* const* keys, const float* vals, uint64_t n) {
  engine_box* b = static_cast<engine_box*>(h);
  try {
    if (b->lof) b->lof->set_row(id, make_sfv(keys, vals, n));
    else b->nn->set_row(id, make_sfv(keys, vals, n));
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}
int oracle_nn_delete_row(void* h, const char* id) {
  return static_cast<engine_box*>(h)->nn->get_table()->delete_row(id) ? 1 : 0;
}
int oracle_nn_query_datum(void* h, const char* const* keys, const float* vals, uint64_t n, uint64_t k,
                          int similar, char* ids_buf, uint64_t ids_cap, float* scores, uint64_t* count) {
  engine_box* b = static_cast<engine_box*>(h);
  std::vector<std::pair<std::string, float> > r;
  sfv_t q = make_sfv(keys, vals, n);
  if (similar) b->nn->similar_row(q, r, k); else b->nn->neighbor_row(q, r, k);
  return emit(r, ids_buf, ids_cap, scores, count);
}
int oracle_nn_query_id(void* h, const char* id, uint64_t k, int similar, char* ids_buf, uint64_t ids_cap,
                       float* scores, uint64_t* count) {
  engine_box* b = static_cast<engine_box*>(h);
  std::vector<std::pair<std::string, float> > r;
  std::string sid(id);
  if (similar) b->nn->similar_row(sid, r, k); else b->nn->neighbor_row(sid, r, k);
  return emit(r, ids_buf, ids_cap, scores, count);
}
int oracle_nn_get_signature(void* h, const char* id, uint64_t* sig, float* norm) {
  engine_box* b = static_cast<engine_box*>(h);
  std::pair<bool, uint64_t> hit = b->nn->get_table()->exact_match(id);
  if (!hit.first) return 0;
  memcpy(sig, b->nn->get_table()->sig(hit.second), b->nn->get_table()->words() * 8);
  if (norm) *norm = b->nn->get_table()->nfloat() ? b->nn->get_table()->get_float(hit.second, 0) : 0.0f;
  return 1;
}
int oracle_nn_get_all_rows(void* h, char* ids_buf, uint64_t ids_cap, uint64_t* count) {
  engine_box* b = static_cast<engine_box*>(h);
  std::vector<std::string> ids = b->nn->get_all_row_ids();
  std::vector<std::pair<std::string, float> > r;
  std::vector<float> dummy(ids.size() + 1);
  for (size_t i = 0; i < ids.size(); ++i) r.push_back(std::make_pair(ids[i], 0.0f));
  return emit(r, ids_buf, ids_cap, dummy.data(), count);
}
// MIX between two oracle engines, push_mixer order (push_mixer.cpp:335-408):
// a pulls from b with a's argument and vice versa, then both push.
void oracle_nn_mix_pair(void* ha, void* hb) {
  nearest_neighbor* a = static_cast<engine_box*>(ha)->nn.get();
  nearest_neighbor* b = static_cast<engine_box*>(hb)->nn.get();
  version_clock arg_a = a->get_mixable()->get_argument();
  version_clock arg_b = b->get_mixable()->get_argument();
  std::vector<packed_row> from_b, from_a;
  b->get_mixable()->pull(arg_a, from_b);
  a->get_mixable()->pull(arg_b, from_a);
  b->get_mixable()->push(from_a);
  a->get_mixable()->push(from_b);
}
float oracle_lof_calc_score_datum(void* h, const char* const* keys, const float* vals, uint64_t n) {
  return static_cast<engine_box*>(h)->lof->calc_anomaly_score(make_sfv(keys, vals, n));
}
float oracle_lof_calc_score_id(void* h, const char* id) {
  return static_cast<engine_box*>(h)->lof->calc_anomaly_score(std::string(id));
}
int oracle_lof_get_params(void* h, const char* id, float* kdist, float* lrd) {
  column_table* t = static_cast<engine_box*>(h)->lof->lof_table();
  std::pair<bool, uint64_t> hit = t->exact_match(id);
  if (!hit.first) return 0;
  *kdist = t->get_float(hit.second, 0); *lrd = t->get_float(hit.second, 1);
  return 1;
}

}  // extern "C"
