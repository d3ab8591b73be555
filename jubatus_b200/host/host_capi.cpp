// Flat C entry points over jubann_host.hpp so the host layer (string ids, datums, JSON
// configs, MIX, recommender, light_lof) can be driven from Python tests through ctypes.
// This is a test/bindings convenience; a C++ server links jubann_host.hpp directly
// (INTEGRATION.md).  Results with string ids come back '\n'-joined.
#include "jubann_host.hpp"

#include <string.h>

#include <sstream>

using namespace jubatus_b200;

namespace {
thread_local std::string g_err;

struct host_box {
  std::shared_ptr<jubann_api> api;
  std::unique_ptr<nearest_neighbor_serv> serv;             // jubanearest_neighbor
  std::shared_ptr<gpu_nearest_neighbor> engine;            // nested engines
  std::shared_ptr<datum_to_fv_converter> converter;
  std::unique_ptr<nearest_neighbor_recommender> recommender;
  std::unique_ptr<light_lof> lof;
  std::unique_ptr<nearest_neighbor_classifier> classifier;   // jubaclassifier, method NN
  gpu_nearest_neighbor* nn() { return serv ? serv->get_driver()->engine() : engine.get(); }
  const datum_to_fv_converter& conv() { return serv ? serv->get_driver()->converter() : *converter; }
  datum_to_fv_converter& conv_mut() { return serv ? serv->get_driver()->converter_mut() : *converter; }
};

datum make_datum(const char* const* skeys, const char* const* svals, uint64_t ns, const char* const* nkeys,
                 const double* nvals, uint64_t nn) {
  datum d;
  for (uint64_t i = 0; i < ns; ++i) d.add_string(skeys[i], svals[i]);
  for (uint64_t i = 0; i < nn; ++i) d.add_number(nkeys[i], nvals[i]);
  return d;
}

int emit(const neighbor_result& r, char* ids_buf, uint64_t ids_cap, float* scores, uint64_t scores_cap, uint64_t* count) {
  std::string joined;
  if (r.size() > scores_cap) { g_err = "scores buffer too small"; return -1; }
  for (size_t i = 0; i < r.size(); ++i) { joined += r[i].first; joined += '\n'; scores[i] = r[i].second; }
  if (joined.size() + 1 > ids_cap) { g_err = "ids buffer too small"; return -1; }
  memcpy(ids_buf, joined.c_str(), joined.size() + 1);
  *count = r.size();
  return 0;
}

template <typename F>
int guarded(F f) {
  try { return f(); }
  catch (const std::exception& e) { g_err = e.what(); return -1; }
}
}  // namespace

extern "C" {

const char* jubahost_last_error() { return g_err.c_str(); }

// kind: "nearest_neighbor" | "recommender" | "anomaly"; config = the server's JSON config
void* jubahost_create_devices(const char* kind, const char* config_json, const char* lib_path, const char* my_id,
                              const int* devices, int ndev);
void* jubahost_create(const char* kind, const char* config_json, const char* lib_path, const char* my_id, int device) {
  const std::vector<int> devs = devices_from_env(device);
  return jubahost_create_devices(kind, config_json, lib_path, my_id, devs.data(), static_cast<int>(devs.size()));
}
// the same with the GPUs the server process may use: more than one -> the table is sharded over
// them inside libjubann.so (jubann_create_multi)
void* jubahost_create_devices(const char* kind, const char* config_json, const char* lib_path, const char* my_id,
                              const int* devices, int ndev) {
  host_box* b = new host_box;
  const std::vector<int> device(devices, devices + (ndev > 0 ? ndev : 0));
  int rc = guarded([&]() {
    if (device.empty()) throw invalid_parameter("device list is empty");
    b->api = jubann_api::load(lib_path ? lib_path : "libjubann.so");
    const std::string k(kind);
    if (k == "nearest_neighbor") {
      b->serv.reset(new nearest_neighbor_serv(b->api, my_id ? my_id : "", device));
      b->serv->set_config(config_json);
      return 0;
    }
    json::ptr root;
    try { root = json::parse(config_json); } catch (const json::parse_error& e) { throw config_exception(e.what()); }
    json::ptr m = root->find("method"), p = root->find("parameter"), c = root->find("converter");
    if (!m || m->kind != json::value::STR) throw config_exception("\"method\" is required");
    if (!c) throw config_exception("\"converter\" is required");
    b->converter = datum_to_fv_converter::from_config(c);
    if (!p || p->kind != json::value::OBJ) throw config_exception("\"parameter\" is required");
    json::ptr im = p->find("method");
    if (!im || im->kind != json::value::STR) throw config_exception("nested \"method\" is required");
    if (k == "recommender") {
      // recommender_serv.cpp:93-119 -> recommender_factory; only the nearest_neighbor_recommender family is on this path
      if (m->str != "nearest_neighbor_recommender") throw unsupported_method(m->str);
      for (size_t i = 0; i < p->obj.size(); ++i) {
        const std::string& key = p->obj[i].first;
        if (key != "method" && key != "parameter" && key != "unlearner" && key != "unlearner_parameter")
          throw config_exception("unknown parameter key: " + key);
      }
      std::shared_ptr<unlearner_base> unl;
      if (json::ptr u = p->find("unlearner")) {
        if (u->kind != json::value::STR) throw config_exception("\"unlearner\" must be a string");
        unl = create_unlearner(u->str, p->find("unlearner_parameter"));
      } else if (p->find("unlearner_parameter")) {
        throw config_exception("\"unlearner_parameter\" without \"unlearner\"");
      }
      b->engine = create_nearest_neighbor(im->str, p->find("parameter"), b->api, my_id ? my_id : "", device);
      b->recommender.reset(new nearest_neighbor_recommender(b->engine, unl));
    } else if (k == "anomaly") {
      // anomaly_serv.cpp:112-142 -> anomaly_factory; light_lof only
      if (m->str != "light_lof") throw unsupported_method(m->str);
      int nn_num = 10, rnn_num = 30;
      for (size_t i = 0; i < p->obj.size(); ++i) {
        const std::string& key = p->obj[i].first;
        const json::ptr& v = p->obj[i].second;
        if (key == "nearest_neighbor_num" || key == "reverse_nearest_neighbor_num") {
          if (v->kind != json::value::NUM || !v->is_integer) throw config_exception("\"" + key + "\" must be an integer");
          (key == "nearest_neighbor_num" ? nn_num : rnn_num) = static_cast<int>(v->num);
        } else if (key == "unlearner" || key == "unlearner_parameter") {
          // handled below
        } else if (key != "method" && key != "parameter") {
          throw config_exception("unknown parameter key: " + key);
        }
      }
      std::shared_ptr<unlearner_base> unl;
      if (json::ptr u = p->find("unlearner")) {
        if (u->kind != json::value::STR) throw config_exception("\"unlearner\" must be a string");
        unl = create_unlearner(u->str, p->find("unlearner_parameter"));
      } else if (p->find("unlearner_parameter")) {
        throw config_exception("\"unlearner_parameter\" without \"unlearner\"");
      }
      b->engine = create_nearest_neighbor(im->str, p->find("parameter"), b->api, my_id ? my_id : "", device);
      b->lof.reset(new light_lof(b->engine, nn_num, rnn_num, unl));
    } else if (k == "classifier") {
      // classifier_serv.cpp -> classifier_factory; only the NN method is on this path (config/classifier/nn.json)
      if (m->str != "NN" && m->str != "nearest_neighbor") throw unsupported_method(m->str);
      int nn_num = -1;
      double alpha = -1.0;
      for (size_t i = 0; i < p->obj.size(); ++i) {
        const std::string& key = p->obj[i].first;
        const json::ptr& v = p->obj[i].second;
        if (key == "nearest_neighbor_num") {
          if (v->kind != json::value::NUM || !v->is_integer) throw config_exception("\"nearest_neighbor_num\" must be an integer");
          nn_num = static_cast<int>(v->num);
        } else if (key == "local_sensitivity") {
          if (v->kind != json::value::NUM) throw config_exception("\"local_sensitivity\" must be a number");
          alpha = v->num;
        } else if (key != "method" && key != "parameter" && key != "unlearner" && key != "unlearner_parameter") {
          throw config_exception("unknown parameter key: " + key);
        }
      }
      if (nn_num < 1) throw config_exception("\"nearest_neighbor_num\" is required and must be >= 1");
      if (alpha < 0) throw config_exception("\"local_sensitivity\" is required and must be >= 0");
      std::shared_ptr<unlearner_base> unl;
      if (json::ptr u = p->find("unlearner")) {
        if (u->kind != json::value::STR) throw config_exception("\"unlearner\" must be a string");
        unl = create_unlearner(u->str, p->find("unlearner_parameter"));
      }
      b->engine = create_nearest_neighbor(im->str, p->find("parameter"), b->api, my_id ? my_id : "", device);
      b->classifier.reset(new nearest_neighbor_classifier(b->engine, static_cast<size_t>(nn_num), static_cast<float>(alpha), unl));
    } else {
      throw jubatus_exception("unknown server kind: " + k);
    }
    return 0;
  });
  if (rc) { delete b; return 0; }
  return b;
}
void jubahost_destroy(void* h) { delete static_cast<host_box*>(h); }

// datum -> sfv only (no device): keys come back '\n'-joined
int jubahost_convert(const char* converter_json, const char* const* skeys, const char* const* svals, uint64_t ns,
                     const char* const* nkeys, const double* nvals, uint64_t nn, char* keys_buf, uint64_t keys_cap,
                     float* vals, uint64_t vals_cap, uint64_t* count) {
  return guarded([&]() {
    json::ptr c;
    try { c = json::parse(converter_json); } catch (const json::parse_error& e) { throw config_exception(e.what()); }
    std::shared_ptr<datum_to_fv_converter> conv = datum_to_fv_converter::from_config(c);
    sfv_t sfv;
    conv->convert(make_datum(skeys, svals, ns, nkeys, nvals, nn), sfv);
    return emit(sfv, keys_buf, keys_cap, vals, vals_cap, count);
  });
}

int jubahost_clear(void* h) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (b->serv) b->serv->clear(); else if (b->recommender) b->recommender->clear();
    else if (b->classifier) b->classifier->clear(); else b->lof->clear();
    if (!b->serv) b->converter->clear_weights();
    return 0;
  });
}
uint64_t jubahost_size(void* h) { return static_cast<host_box*>(h)->nn()->get_table()->size(); }

// set_row (nearest_neighbor) / update_row (recommender) / add-with-id (anomaly: light_lof::set_row)
int jubahost_set_row(void* h, const char* id, const char* const* skeys, const char* const* svals, uint64_t ns,
                     const char* const* nkeys, const double* nvals, uint64_t nn) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    datum d = make_datum(skeys, svals, ns, nkeys, nvals, nn);
    if (b->serv) { b->serv->set_row(id, d); return 0; }
    sfv_t v;
    b->conv_mut().convert_and_update_weight(d, v);   // a stored row moves the global weights
    if (b->recommender) b->recommender->update_row(id, v); else b->lof->set_row(id, v);
    return 0;
  });
}
// batched set_row over n datums that carry only number values: nkeys/nvals are flattened, offs[n+1]
int jubahost_set_rows(void* h, uint64_t n, const char* const* ids, const uint64_t* offs, const char* const* nkeys,
                      const double* nvals) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->serv && !b->lof) throw jubatus_exception("set_rows: nearest_neighbor and anomaly servers only");
    std::vector<std::string> vid(n);
    std::vector<datum> ds(n);
    for (uint64_t i = 0; i < n; ++i) {
      vid[i] = ids[i];
      for (uint64_t j = offs[i]; j < offs[i + 1]; ++j) ds[i].add_number(nkeys[j], nvals[j]);
    }
    if (b->lof) {   // anomaly: bulk load (light_lof::set_rows_bulk)
      std::vector<sfv_t> v(n);
      for (uint64_t i = 0; i < n; ++i) b->conv_mut().convert_and_update_weight(ds[i], v[i]);
      b->lof->set_rows_bulk(vid, v);
      return 0;
    }
    b->serv->get_driver()->set_rows(vid, ds);
    return 0;
  });
}
int jubahost_delete_row(void* h, const char* id) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (b->recommender) { b->recommender->clear_row(id); return 0; }
    if (b->lof) return b->lof->delete_row(id) ? 0 : 1;   // both of its tables
    return b->nn()->delete_row(id) ? 0 : 1;
  });
}
int jubahost_query_datum(void* h, const char* const* skeys, const char* const* svals, uint64_t ns,
                         const char* const* nkeys, const double* nvals, uint64_t nn, uint64_t k, int similar,
                         char* ids_buf, uint64_t ids_cap, float* scores, uint64_t scores_cap, uint64_t* count) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    datum d = make_datum(skeys, svals, ns, nkeys, nvals, nn);
    neighbor_result r;
    if (b->serv) {
      r = similar ? b->serv->similar_row_from_datum(d, k) : b->serv->neighbor_row_from_datum(d, k);
    } else {
      sfv_t v;
      b->conv().convert(d, v);
      if (b->recommender) { if (similar) b->recommender->similar_row(v, r, k); else b->recommender->neighbor_row(v, r, k); }
      else { if (similar) b->nn()->similar_row(v, r, k); else b->nn()->neighbor_row(v, r, k); }
    }
    return emit(r, ids_buf, ids_cap, scores, scores_cap, count);
  });
}
int jubahost_query_id(void* h, const char* id, uint64_t k, int similar, char* ids_buf, uint64_t ids_cap, float* scores,
                      uint64_t scores_cap, uint64_t* count) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    neighbor_result r;
    if (b->serv) r = similar ? b->serv->similar_row_from_id(id, k) : b->serv->neighbor_row_from_id(id, k);
    else if (b->recommender) { if (similar) b->recommender->similar_row(std::string(id), r, k); else b->recommender->neighbor_row(std::string(id), r, k); }
    else { if (similar) b->nn()->similar_row(std::string(id), r, k); else b->nn()->neighbor_row(std::string(id), r, k); }
    return emit(r, ids_buf, ids_cap, scores, scores_cap, count);
  });
}
int jubahost_get_all_rows(void* h, char* ids_buf, uint64_t ids_cap, uint64_t* count) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    std::vector<std::string> ids;
    b->nn()->get_all_row_ids(ids);
    neighbor_result r;
    for (size_t i = 0; i < ids.size(); ++i) r.push_back(std::make_pair(ids[i], 0.0f));
    std::vector<float> dummy(ids.size() + 1);
    return emit(r, ids_buf, ids_cap, dummy.data(), dummy.size(), count);
  });
}
int jubahost_get_signature(void* h, const char* id, uint64_t* sig, float* norm) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    std::pair<bool, uint64_t> hit = b->nn()->get_table()->exact_match(id);
    if (!hit.first) return 1;
    packed_row r = b->nn()->get_table()->get_row(hit.second);
    memcpy(sig, r.sig.data(), r.sig.size() * 8);
    if (norm) *norm = r.norm;
    return 0;
  });
}
// push_mixer order (framework/mixer/push_mixer.cpp:335-408) between two in-process engines
int jubahost_mix_pair(void* ha, void* hb) {
  host_box* a = static_cast<host_box*>(ha);
  host_box* b = static_cast<host_box*>(hb);
  return guarded([&]() {
    mixable_versioned_table* ma = a->nn()->get_mixable();
    mixable_versioned_table* mb = b->nn()->get_mixable();
    version_clock arg_a = ma->get_argument(), arg_b = mb->get_argument();
    std::vector<packed_row> from_b, from_a;
    mb->pull(arg_a, from_b);
    ma->pull(arg_b, from_a);
    mb->push(from_a);
    ma->push(from_b);
    // the converters' global weights are the second mixable of the driver: every replica's
    // count diff is summed and the sum goes into every master (linear_mixer.cpp:547-561,613-628)
    weight_manager::counts da, db, acc;
    a->conv_mut().weights().get_diff(da);
    b->conv_mut().weights().get_diff(db);
    weight_manager::mix(da, acc);
    weight_manager::mix(db, acc);
    a->conv_mut().weights().put_diff(acc);
    b->conv_mut().weights().put_diff(acc);
    return 0;
  });
}
// the id rank 0 ships to the other replicas / shards (JUBANN_COMM_ID_BYTES bytes)
int jubahost_comm_unique_id(void* h, void* id) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() { b->api->check(b->api->comm_unique_id(id)); return 0; });
}
// joins this server's table to the communicator of the replicas it mixes with
int jubahost_comm_init(void* h, const void* id, uint32_t rank, uint32_t world) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    b->api->check(b->api->comm_init(b->nn()->get_table()->handle(), id, rank, world));
    return 0;
  });
}
// one MIX round with every replica of the communicator (a collective: all of them call it);
// stats: rows sent, rows gathered, rows stored, bytes of this replica's blob
int jubahost_mix_allgather(void* h, uint64_t* stats) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    mixable_versioned_table::mix_stats st;
    if (b->serv) {
      st = b->serv->get_driver()->mix_allgather();
    } else {   // recommender / anomaly / classifier: the engine's table and this box's converter
      weight_manager::counts mine;
      b->conv_mut().weights().get_diff(mine);
      // extra = [u64 length | converter weight counts | light_lof (kdist, lrd) diff]
      const std::string counts = weight_manager::pack_counts(mine);
      std::string extra;
      const uint64_t clen = counts.size();
      extra.append(reinterpret_cast<const char*>(&clen), 8);
      extra += counts;
      if (b->lof) extra += b->lof->pack_param_diff();
      std::vector<std::string> extras;
      st = b->nn()->mix_allgather(extra, &extras);
      weight_manager::counts acc;
      for (size_t r = 0; r < extras.size(); ++r) {
        if (extras[r].size() < 8) continue;
        uint64_t len; memcpy(&len, extras[r].data(), 8);
        if (extras[r].size() < 8 + len) throw jubatus_exception("mix: truncated blob");
        if (len) weight_manager::mix(weight_manager::unpack_counts(extras[r].substr(8, len)), acc);
        if (b->lof) b->lof->apply_param_diff(extras[r], 8 + len);   // rank order: the same result on every replica
      }
      b->conv_mut().weights().put_diff(acc);
      if (b->lof) b->lof->clear_param_dirty();
    }
    if (stats) { stats[0] = st.rows_sent; stats[1] = st.rows_gathered; stats[2] = st.rows_stored; stats[3] = st.blob_bytes; }
    return 0;
  });
}
// driver->pack(): returns the blob size; copies it when the buffer is large enough
int64_t jubahost_pack(void* h, char* buf, uint64_t cap) {
  host_box* b = static_cast<host_box*>(h);
  int64_t n = -1;
  int rc = guarded([&]() {
    std::string blob = b->serv ? b->serv->get_driver()->pack() : b->nn()->pack();   // the server's pack carries the converter weights too
    if (blob.size() <= cap) memcpy(buf, blob.data(), blob.size());
    n = static_cast<int64_t>(blob.size());
    return 0;
  });
  return rc ? -1 : n;
}
int jubahost_unpack(void* h, const char* buf, uint64_t n) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (b->serv) b->serv->get_driver()->unpack(std::string(buf, n)); else b->nn()->unpack(std::string(buf, n));
    return 0;
  });
}
// ---- model files (nearest_neighbor kind): server_base::save / load / load_file ----
uint32_t jubahost_crc32(const char* data, uint64_t n, uint32_t crc) { return model_file::calc_crc32(data, n, crc); }
int jubahost_set_storage(void* h, const char* datadir, const char* eth, int port) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->serv) throw jubatus_exception("model files are served by the nearest_neighbor kind");
    b->serv->set_storage(datadir, eth, port);
    return 0;
  });
}
int jubahost_save(void* h, const char* id, char* path_buf, uint64_t cap) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->serv) throw jubatus_exception("model files are served by the nearest_neighbor kind");
    std::map<std::string, std::string> r = b->serv->save(id);
    const std::string out = r.begin()->first + "\n" + r.begin()->second;
    if (out.size() + 1 > cap) throw jubatus_exception("path buffer too small");
    memcpy(path_buf, out.c_str(), out.size() + 1);
    return 0;
  });
}
int jubahost_load(void* h, const char* id) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->serv) throw jubatus_exception("model files are served by the nearest_neighbor kind");
    b->serv->load(id);
    return 0;
  });
}
int jubahost_load_file(void* h, const char* path) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->serv) throw jubatus_exception("model files are served by the nearest_neighbor kind");
    b->serv->load_file(path);
    return 0;
  });
}
int jubahost_get_config(void* h, char* buf, uint64_t cap) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->serv) throw jubatus_exception("get_config is served by the nearest_neighbor kind");
    const std::string c = b->serv->get_config();
    if (c.size() + 1 > cap) throw jubatus_exception("config buffer too small");
    memcpy(buf, c.c_str(), c.size() + 1);
    return 0;
  });
}
// container only, no engine: write / check a model file around an arbitrary payload (CPU tests)
int64_t jubahost_model_write(const char* type, const char* id, const char* config, uint64_t user_data_version,
                             const char* payload, uint64_t payload_n, uint64_t timestamp, char* out, uint64_t cap) {
  int64_t n = -1;
  int rc = guarded([&]() {
    std::ostringstream os;
    model_file::save(os, type, id, config, user_data_version, std::string(payload, payload_n), timestamp);
    const std::string f = os.str();
    if (f.size() <= cap) memcpy(out, f.data(), f.size());
    n = static_cast<int64_t>(f.size());
    return 0;
  });
  return rc ? -1 : n;
}
int64_t jubahost_model_read(const char* file, uint64_t file_n, const char* expected_type, const char* expected_config,
                            int overwrite_config, uint64_t user_data_version, char* payload, uint64_t cap) {
  int64_t n = -1;
  int rc = guarded([&]() {
    std::istringstream is(std::string(file, file_n));
    model_file::system_data sd;
    const std::string p = model_file::load(is, expected_type, expected_config, overwrite_config != 0, user_data_version, &sd);
    if (p.size() <= cap) memcpy(payload, p.data(), p.size());
    n = static_cast<int64_t>(p.size());
    return 0;
  });
  return rc ? -1 : n;
}
// converter with global weights (CPU tests): `n_update` string-valued datums (one string value each,
// key skeys[i] / value svals[i]) are converted with convert_and_update_weight, then the last
// (skeys[n_update], svals[n_update]) with convert; the latter's features come back.
int jubahost_convert_weighted(const char* converter_json, const char* const* skeys, const char* const* svals,
                              uint64_t n_update, char* keys_buf, uint64_t keys_cap, float* vals, uint64_t vals_cap,
                              uint64_t* count) {
  return guarded([&]() {
    json::ptr root;
    try { root = json::parse(converter_json); } catch (const json::parse_error& e) { throw config_exception(e.what()); }
    std::shared_ptr<datum_to_fv_converter> c = datum_to_fv_converter::from_config(root);
    sfv_t v;
    for (uint64_t i = 0; i < n_update; ++i) { datum d; d.add_string(skeys[i], svals[i]); c->convert_and_update_weight(d, v); }
    datum q; q.add_string(skeys[n_update], svals[n_update]);
    c->convert(q, v);
    std::string joined;
    if (v.size() > vals_cap) throw jubatus_exception("values buffer too small");
    for (size_t i = 0; i < v.size(); ++i) { joined += v[i].first; joined += '\n'; vals[i] = v[i].second; }
    if (joined.size() + 1 > keys_cap) throw jubatus_exception("keys buffer too small");
    memcpy(keys_buf, joined.c_str(), joined.size() + 1);
    *count = v.size();
    return 0;
  });
}
// unlearner policy alone (CPU tests): touch ids in order, report "<position>:<victim>" per eviction
int jubahost_unlearner_trace(const char* name, const char* param_json, const char* const* ids, uint64_t n, char* out,
                             uint64_t cap, uint64_t* final_size) {
  return guarded([&]() {
    json::ptr param;
    try { param = json::parse(param_json); } catch (const json::parse_error& e) { throw config_exception(e.what()); }
    std::shared_ptr<unlearner_base> u = create_unlearner(name, param);
    std::string log;
    uint64_t pos = 0;
    u->set_callback([&](const std::string& victim) { log += std::to_string(pos) + ":" + victim + "\n"; });
    for (pos = 0; pos < n; ++pos) {
      const std::string id(ids[pos]);
      if (!id.empty() && id[0] == '-') u->remove(id.substr(1));   // "-x" removes x
      else u->touch(id);
    }
    if (log.size() + 1 > cap) throw jubatus_exception("trace buffer too small");
    memcpy(out, log.c_str(), log.size() + 1);
    *final_size = u->size();
    return 0;
  });
}
// ---- jubaclassifier, method NN ----
int jubahost_train(void* h, const char* label, const char* const* skeys, const char* const* svals, uint64_t ns,
                   const char* const* nkeys, const double* nvals, uint64_t nn) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->classifier) throw jubatus_exception("train: classifier server only");
    sfv_t v;
    b->conv_mut().convert_and_update_weight(make_datum(skeys, svals, ns, nkeys, nvals, nn), v);
    b->classifier->train(v, label);
    return 0;
  });
}
// classify over n number-valued datums (flattened like jubahost_set_rows); the result is one line
// per datum: "label\tscore\tlabel\tscore..."
int jubahost_classify(void* h, uint64_t n, const uint64_t* offs, const char* const* nkeys, const double* nvals, char* out,
                      uint64_t cap) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->classifier) throw jubatus_exception("classify: classifier server only");
    std::vector<sfv_t> v(n);
    for (uint64_t i = 0; i < n; ++i) {
      datum d;
      for (uint64_t j = offs[i]; j < offs[i + 1]; ++j) d.add_number(nkeys[j], nvals[j]);
      b->conv().convert(d, v[i]);
    }
    std::vector<nearest_neighbor_classifier::classify_result> res;
    b->classifier->classify_batch(v, res);
    std::string s;
    char num[48];
    for (size_t i = 0; i < res.size(); ++i) {
      for (size_t j = 0; j < res[i].size(); ++j) {
        snprintf(num, sizeof(num), "%.9g", static_cast<double>(res[i][j].second));
        if (j) s += '\t';
        s += res[i][j].first; s += '\t'; s += num;
      }
      s += '\n';
    }
    if (s.size() + 1 > cap) throw jubatus_exception("classify buffer too small");
    memcpy(out, s.c_str(), s.size() + 1);
    return 0;
  });
}
int jubahost_get_labels(void* h, char* out, uint64_t cap) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->classifier) throw jubatus_exception("get_labels: classifier server only");
    std::string s;
    const std::map<std::string, uint64_t> l = b->classifier->get_labels();
    for (std::map<std::string, uint64_t>::const_iterator it = l.begin(); it != l.end(); ++it)
      s += it->first + "\t" + std::to_string(it->second) + "\n";
    if (s.size() + 1 > cap) throw jubatus_exception("labels buffer too small");
    memcpy(out, s.c_str(), s.size() + 1);
    return 0;
  });
}
int jubahost_set_label(void* h, const char* label) {
  host_box* b = static_cast<host_box*>(h);
  int r = 0;
  int rc = guarded([&]() {
    if (!b->classifier) throw jubatus_exception("set_label: classifier server only");
    r = b->classifier->set_label(label) ? 1 : 0;
    return 0;
  });
  return rc ? -1 : r;
}
int jubahost_delete_label(void* h, const char* label) {
  host_box* b = static_cast<host_box*>(h);
  int r = 0;
  int rc = guarded([&]() {
    if (!b->classifier) throw jubatus_exception("delete_label: classifier server only");
    r = b->classifier->delete_label(label) ? 1 : 0;
    return 0;
  });
  return rc ? -1 : r;
}
int jubahost_dump_json(void* h, char* buf, uint64_t cap) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    std::string j = b->nn()->get_table()->dump_json();
    if (j.size() + 1 > cap) throw jubatus_exception("buffer too small");
    memcpy(buf, j.c_str(), j.size() + 1);
    return 0;
  });
}
int jubahost_enable_batching(void* h, uint64_t max_batch, uint64_t max_wait_us) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->serv) throw jubatus_exception("batching: nearest_neighbor server only");
    b->serv->enable_batching(max_batch, max_wait_us);
    return 0;
  });
}
int jubahost_batch_stats(void* h, uint64_t* batches, uint64_t* queries) {
  host_box* b = static_cast<host_box*>(h);
  *batches = b->serv ? b->serv->batches_run() : 0;
  *queries = b->serv ? b->serv->queries_batched() : 0;
  return 0;
}
int jubahost_calc_score_datum(void* h, const char* const* nkeys, const double* nvals, uint64_t nn, float* score) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->lof) throw jubatus_exception("calc_score: anomaly server only");
    sfv_t v;
    b->conv().convert(make_datum(0, 0, 0, nkeys, nvals, nn), v);
    *score = b->lof->calc_anomaly_score(v);
    return 0;
  });
}
int jubahost_calc_score_id(void* h, const char* id, float* score) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->lof) throw jubatus_exception("calc_score: anomaly server only");
    *score = b->lof->calc_anomaly_score(std::string(id));
    return 0;
  });
}
int jubahost_lof_params(void* h, const char* id, float* kdist, float* lrd) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() { return b->lof && b->lof->get_row_parameter(id, kdist, lrd) ? 0 : 1; });
}

// ---- number-only datums whose keys are decimal feature ids (the synthetic workload of SURVEY 8d):
// the datum is built here, in C++, and takes the same road as any other (converter -> sfv -> engine);
// what is saved is the Python-side marshalling of 100 key strings per datum. ----
static datum numeric_datum(uint32_t nnz, const uint32_t* fids, const float* vals) {
  datum d;
  d.num_values.reserve(nnz);
  char key[16];
  for (uint32_t j = 0; j < nnz; ++j) {
    char* e = key + sizeof key;   // decimal digits, written backwards
    char* p = e;
    uint32_t f = fids[j];
    do { *--p = static_cast<char>('0' + f % 10); f /= 10; } while (f);
    d.num_values.push_back(std::make_pair(std::string(p, e - p), static_cast<double>(vals[j])));
  }
  return d;
}
// rows "<prefix><first_id + i>": batched set_rows (nearest_neighbor) or light_lof::set_rows_bulk (anomaly)
int jubahost_set_rows_numeric(void* h, uint64_t n, const char* prefix, uint64_t first_id, uint32_t nnz, const uint32_t* fids,
                              const float* vals) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->serv && !b->lof) throw jubatus_exception("set_rows: nearest_neighbor and anomaly servers only");
    std::vector<std::string> vid(n);
    for (uint64_t i = 0; i < n; ++i) vid[i] = std::string(prefix ? prefix : "") + std::to_string(first_id + i);
    std::vector<sfv_t> v;
    const unsigned threads = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    b->conv_mut().convert_and_update_weight_batch(
        n, [&](size_t i) { return numeric_datum(nnz, fids + i * nnz, vals + i * nnz); }, v, threads);
    if (b->lof) b->lof->set_rows_bulk(vid, v); else b->nn()->set_rows(vid, v);
    return 0;
  });
}
// anomaly add(datum) with a given id: set_row, then the new row's score (anomaly_serv.cpp:157-176)
int jubahost_add_numeric(void* h, const char* id, uint32_t nnz, const uint32_t* fids, const float* vals, float* score) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->lof) throw jubatus_exception("add: anomaly server only");
    sfv_t v;
    b->conv_mut().convert_and_update_weight(numeric_datum(nnz, fids, vals), v);
    *score = b->lof->set_row_and_score(id, v);
    return 0;
  });
}
int jubahost_calc_score_numeric(void* h, uint32_t nnz, const uint32_t* fids, const float* vals, float* score) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    if (!b->lof) throw jubatus_exception("calc_score: anomaly server only");
    sfv_t v;
    b->conv().convert(numeric_datum(nnz, fids, vals), v);
    *score = b->lof->calc_anomaly_score(v);
    return 0;
  });
}
// out[0]: order-independent digest of (key, signature, norm) over the whole table; out[1]: of the
// light_lof (kdist, lrd) table (0 for other servers).  Replicas that hold the same rows agree.
int jubahost_table_digest(void* h, uint64_t* out) {
  host_box* b = static_cast<host_box*>(h);
  return guarded([&]() {
    const std::shared_ptr<device_table> t = b->nn()->get_table();
    const uint64_t n = t->size();
    uint64_t acc = 0;
    const uint64_t step = 1u << 18;
    std::vector<uint64_t> idx;
    std::vector<packed_row> rows;
    for (uint64_t r0 = 0; r0 < n; r0 += step) {
      const uint64_t r1 = std::min(n, r0 + step);
      idx.resize(r1 - r0);
      for (uint64_t i = r0; i < r1; ++i) idx[i - r0] = i;
      t->get_rows(idx, rows);
      for (size_t i = 0; i < rows.size(); ++i) {
        uint64_t x = calc_string_hash(rows[i].key);
        for (size_t w = 0; w < rows[i].sig.size(); ++w) { x ^= rows[i].sig[w]; x *= 0x9E3779B97F4A7C15ULL; x ^= x >> 29; }
        uint32_t nb; memcpy(&nb, &rows[i].norm, 4);
        x ^= nb; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33;
        acc += x;
      }
    }
    out[0] = acc;
    out[1] = b->lof ? b->lof->param_digest() : 0;
    return 0;
  });
}

}  // extern "C"
