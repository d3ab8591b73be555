// A C++ host with no Python in the process: the table of one nearest_neighbor engine sharded over
// several GPUs behind the C-ABI (jubann_create_multi: what nearest_neighbor_serv.cpp:98-109 becomes
// on a multi-GPU box), checked against the single-GPU handle and against a brute-force ranking
// computed here -- ids, order and score bit patterns must be identical.  Also drives the same table
// through the server facade (nearest_neighbor_serv with a device list).
//
//   multi_gpu_test <path to libjubann.so> <ndev>
//
// tests/test_cpp_multi_gpu.py builds and runs it (ndev = the GPUs of the box).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "jubann_host.hpp"

namespace {

uint64_t splitmix(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

int failures = 0;
#define EXPECT(cond, ...)                         \
  do {                                            \
    if (!(cond)) {                                \
      ++failures;                                 \
      std::fprintf(stderr, "FAILED %s:%d: ", __FILE__, __LINE__); \
      std::fprintf(stderr, __VA_ARGS__);          \
      std::fprintf(stderr, "\n");                 \
    }                                             \
  } while (0)

struct answer {
  std::vector<uint64_t> idx;
  std::vector<float> score;
  std::vector<uint32_t> count;
};

answer ask(const jubatus_b200::jubann_api& api, jubann_t* h, const std::vector<uint64_t>& q, const std::vector<float>& qn,
           uint64_t nq, uint32_t k, uint32_t flags) {
  answer a;
  a.idx.assign(nq * k, 0); a.score.assign(nq * k, 0.0f); a.count.assign(nq, 0);
  api.check(api.query_sig(h, nq, q.data(), qn.empty() ? 0 : qn.data(), k, flags, a.idx.data(), a.score.data(), a.count.data()));
  return a;
}

bool same(const answer& a, const answer& b, uint64_t nq, uint32_t k) {
  for (uint64_t q = 0; q < nq; ++q) {
    if (a.count[q] != b.count[q]) return false;
    for (uint32_t j = 0; j < a.count[q]; ++j) {
      if (a.idx[q * k + j] != b.idx[q * k + j]) return false;
      if (std::memcmp(&a.score[q * k + j], &b.score[q * k + j], 4) != 0) return false;
    }
  }
  return true;
}

// ranking_hamming_bit_vectors restated in a few lines: the k smallest (distance, row index)
answer brute_force_hamming(const std::vector<uint64_t>& rows, uint64_t n, uint32_t w64, uint32_t hash_num,
                           const std::vector<uint64_t>& q, uint64_t nq, uint32_t k) {
  answer a;
  a.idx.assign(nq * k, 0); a.score.assign(nq * k, 0.0f); a.count.assign(nq, 0);
  std::vector<std::pair<uint32_t, uint64_t> > d(n);
  for (uint64_t qi = 0; qi < nq; ++qi) {
    for (uint64_t r = 0; r < n; ++r) {
      uint32_t dist = 0;
      for (uint32_t w = 0; w < w64; ++w) dist += (uint32_t)__builtin_popcountll(rows[r * w64 + w] ^ q[qi * w64 + w]);
      d[r] = std::make_pair(dist, r);
    }
    const uint64_t m = std::min<uint64_t>(k, n);
    std::partial_sort(d.begin(), d.begin() + m, d.end());
    a.count[qi] = (uint32_t)m;
    for (uint64_t j = 0; j < m; ++j) {
      a.idx[qi * k + j] = d[j].second;
      a.score[qi * k + j] = (float)d[j].first / (float)hash_num;
    }
  }
  return a;
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 3) { std::fprintf(stderr, "usage: %s <libjubann.so> <ndev>\n", argv[0]); return 2; }
  const int ndev = std::atoi(argv[2]);
  try {
    std::shared_ptr<jubatus_b200::jubann_api> api = jubatus_b200::jubann_api::load(argv[1]);
    std::vector<int> devices;
    for (int d = 0; d < ndev; ++d) devices.push_back(d);
    uint64_t seed = 0x4A554241ull;

    // ---- C-ABI: one table over `ndev` GPUs against the same table on GPU 0 ----
    const char* methods[] = {"lsh", "euclid_lsh", "minhash"};
    const uint32_t hash_nums[] = {512, 256, 128};
    for (int m = 0; m < 3; ++m) {
      const uint32_t hash_num = hash_nums[m], w64 = (hash_num + 63) / 64, k = 100;
      const uint64_t n = 200000 + 17 * (uint64_t)m, nq = 33;
      const bool euclid = m == 1;
      std::vector<uint64_t> rows(n * w64), q(nq * w64), index(n);
      std::vector<float> norms, qn;
      // few distinct rows => long runs of equal distances whose order only the row index decides
      std::vector<uint64_t> base(64 * w64);
      for (size_t i = 0; i < base.size(); ++i) base[i] = splitmix(seed);
      for (uint64_t r = 0; r < n; ++r) {
        index[r] = r;
        const uint64_t b = splitmix(seed) % 64;
        for (uint32_t w = 0; w < w64; ++w) rows[r * w64 + w] = (r % 3 == 0) ? splitmix(seed) : base[b * w64 + w];
      }
      for (uint64_t i = 0; i < nq; ++i)
        for (uint32_t w = 0; w < w64; ++w) q[i * w64 + w] = (i < 8) ? base[i * w64 + w] : splitmix(seed);
      if (euclid) {
        norms.resize(n); qn.resize(nq);
        for (uint64_t r = 0; r < n; ++r) norms[r] = 0.25f * (float)(splitmix(seed) % 16) + 0.5f;   // ties in the score as well
        for (uint64_t i = 0; i < nq; ++i) qn[i] = 0.5f * (float)(splitmix(seed) % 8) + 1.0f;
      }
      jubann_t *multi = 0, *one = 0;
      api->check(api->create_multi(methods[m], hash_num, devices.data(), ndev, &multi));
      api->check(api->create(methods[m], hash_num, 0, 0, 1, &one));
      // rows arrive in three uneven batches
      const uint64_t cuts[4] = {0, n / 3, n / 3 + 1, n};
      for (int c = 0; c < 3; ++c) {
        const uint64_t a = cuts[c], b = cuts[c + 1];
        api->check(api->set_rows_sig(multi, b - a, index.data() + a, rows.data() + a * w64, euclid ? norms.data() + a : 0));
        api->check(api->set_rows_sig(one, b - a, index.data() + a, rows.data() + a * w64, euclid ? norms.data() + a : 0));
      }
      uint64_t g = 0, l = 0;
      api->check(api->size(multi, &g, &l));
      EXPECT(g == n, "%s: size %llu != %llu", methods[m], (unsigned long long)g, (unsigned long long)n);
      uint32_t rank = 9, world = 0;
      api->check(api->comm_info(multi, &rank, &world));
      EXPECT(world == (uint32_t)ndev, "world %u != %d", world, ndev);
      for (uint32_t flags = 0; flags < 2; ++flags) {
        const answer am = ask(*api, multi, q, qn, nq, k, flags), ao = ask(*api, one, q, qn, nq, k, flags);
        EXPECT(same(am, ao, nq, k), "%s flags=%u: %d GPUs and one GPU disagree", methods[m], flags, ndev);
        if (!euclid && flags == 0) {
          const answer bf = brute_force_hamming(rows, n, w64, hash_num, q, nq, k);
          EXPECT(same(am, bf, nq, k), "%s: the sharded table and the brute-force ranking disagree", methods[m]);
        }
      }
      // neighbor_row(id) on rows of different shards, then delete_row (swap with the last row)
      std::vector<uint64_t> ids;
      for (int d = 0; d < ndev + 1; ++d) ids.push_back((uint64_t)d);
      answer rm, ro;
      rm.idx.assign(ids.size() * k, 0); rm.score.assign(ids.size() * k, 0); rm.count.assign(ids.size(), 0);
      ro = rm;
      api->check(api->query_rows(multi, ids.size(), ids.data(), k, 0, rm.idx.data(), rm.score.data(), rm.count.data()));
      api->check(api->query_rows(one, ids.size(), ids.data(), k, 0, ro.idx.data(), ro.score.data(), ro.count.data()));
      EXPECT(same(rm, ro, ids.size(), k), "%s: query_rows differs", methods[m]);
      api->check(api->move_row(multi, n - 1, 1)); api->check(api->truncate(multi, n - 1));
      api->check(api->move_row(one, n - 1, 1)); api->check(api->truncate(one, n - 1));
      EXPECT(same(ask(*api, multi, q, qn, nq, k, 0), ask(*api, one, q, qn, nq, k, 0), nq, k), "%s: after delete_row", methods[m]);
      api->check(api->destroy(multi));
      api->check(api->destroy(one));
      std::printf("%s hash_num=%u rows=%llu over %d GPU(s): identical to one GPU%s\n", methods[m], hash_num,
                  (unsigned long long)n, ndev, euclid ? "" : " and to the brute-force ranking");
    }

    // ---- the server facade with a device list: the RPC surface unchanged ----
    {
      const std::string config =
          "{\"method\": \"lsh\", \"parameter\": {\"hash_num\": 128},"
          " \"converter\": {\"num_rules\": [{\"key\": \"*\", \"type\": \"num\"}]}}";
      jubatus_b200::nearest_neighbor_serv many(api, "127.0.0.1_9199", devices), single(api, "127.0.0.1_9199", 0);
      many.set_config(config); single.set_config(config);
      for (int r = 0; r < 500; ++r) {
        jubatus_b200::datum d;
        for (int f = 0; f < 20; ++f) {
          char key[32];
          std::snprintf(key, sizeof key, "f%llu", (unsigned long long)(splitmix(seed) % 300));
          d.num_values.push_back(std::make_pair(std::string(key), 1.0 + (double)(splitmix(seed) % 1000) / 1000.0));
        }
        char id[32];
        std::snprintf(id, sizeof id, "row%d", r % 450);      // some ids repeat: updates in place
        many.set_row(id, d); single.set_row(id, d);
      }
      for (int r = 0; r < 20; ++r) {
        char id[32];
        std::snprintf(id, sizeof id, "row%d", r * 7);
        const jubatus_b200::neighbor_result a = many.neighbor_row_from_id(id, 15), b = single.neighbor_row_from_id(id, 15);
        bool eq = a.size() == b.size();
        for (size_t i = 0; eq && i < a.size(); ++i) eq = a[i].first == b[i].first && std::memcmp(&a[i].second, &b[i].second, 4) == 0;
        EXPECT(eq, "nearest_neighbor_serv over %d GPUs: neighbor_row_from_id(%s) differs", ndev, id);
      }
      EXPECT(many.get_all_rows().size() == 450, "get_all_rows");
      std::printf("nearest_neighbor_serv over %d GPU(s): same answers as on one\n", ndev);
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "exception: %s\n", e.what());
    return 1;
  }
  if (failures) { std::fprintf(stderr, "%d check(s) failed\n", failures); return 1; }
  std::printf("ok\n");
  return 0;
}
