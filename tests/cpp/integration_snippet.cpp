// The reference-side binding shown in INTEGRATION.md section 2, as a translation unit: it has to
// keep compiling against jubann_host.hpp (tests/test_integration_snippet_cpu.py runs g++
// -fsyntax-only on it).  Nothing here is executed.
#include <memory>
#include <string>
#include <vector>

#include "jubann_host.hpp"

namespace {

using jubatus_b200::datum;
using jubatus_b200::neighbor_result;
using jubatus_b200::sfv_t;

// nearest_neighbor_serv.cpp:98-109 -> the engine behind the same factory call
std::shared_ptr<jubatus_b200::gpu_nearest_neighbor> make_engine(const std::string& method, const std::string& param_json,
                                                               const std::string& my_id) {
  static std::shared_ptr<jubatus_b200::jubann_api> api = jubatus_b200::jubann_api::load("libjubann.so");
  return jubatus_b200::create_nearest_neighbor(method, jubatus_b200::json::parse(param_json), api, my_id, /*device=*/0);
}

void use_like_nearest_neighbor_base(jubatus_b200::gpu_nearest_neighbor& nn, const sfv_t& fv) {
  neighbor_result ids;
  nn.set_row("id", fv);
  nn.neighbor_row(fv, ids, 10);
  nn.neighbor_row(std::string("id"), ids, 10);
  nn.similar_row(fv, ids, 10);
  std::vector<std::string> all;
  nn.get_all_row_ids(all);
  (void)nn.type();
  (void)nn.get_table();
  (void)nn.get_mixable();
  nn.clear();
}

// recommender_serv.cpp:111-116, anomaly_serv.cpp:134-138, classifier method NN
void embedders(const std::shared_ptr<jubatus_b200::gpu_nearest_neighbor>& nn, const sfv_t& fv) {
  std::shared_ptr<jubatus_b200::unlearner_base> unl =
      jubatus_b200::create_unlearner("lru", jubatus_b200::json::parse("{\"max_size\": 4096}"));
  jubatus_b200::nearest_neighbor_recommender rec(nn, unl);
  rec.update_row("u", fv);
  jubatus_b200::light_lof lof(nn, 10, 30);
  lof.set_row("p", fv);
  (void)lof.calc_anomaly_score(fv);
  jubatus_b200::nearest_neighbor_classifier clf(nn, 128, 1.0f);
  clf.train(fv, "label");
  jubatus_b200::nearest_neighbor_classifier::classify_result scores;
  clf.classify_with_scores(fv, scores);
}

// the whole server facade with the RPC names of nearest_neighbor.idl:6-27 and save / load
void server(const std::string& config, const datum& d) {
  jubatus_b200::nearest_neighbor_serv serv(jubatus_b200::jubann_api::load(), "127.0.0.1_9199", 0);
  serv.set_config(config);
  serv.set_row("id", d);
  (void)serv.neighbor_row_from_id("id", 10);
  (void)serv.neighbor_row_from_datum(d, 10);
  (void)serv.similar_row_from_id("id", 10);
  (void)serv.similar_row_from_datum(d, 10);
  (void)serv.get_all_rows();
  serv.enable_batching(64, 200);
  serv.set_storage("/tmp", "127.0.0.1", 9199);
  (void)serv.save("snapshot");
  (void)serv.load("snapshot");
  serv.clear();
}

}  // namespace

int main() {
  (void)&make_engine; (void)&use_like_nearest_neighbor_base; (void)&embedders; (void)&server;
  return 0;
}
