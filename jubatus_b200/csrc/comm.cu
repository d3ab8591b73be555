// More than one GPU behind the C-ABI: NCCL (over NVLink / NVSwitch) under libjubann.so.
//
// In the reference everything that touches the engine is C++: the engine is created at
// jubatus/server/server/nearest_neighbor_serv.cpp:98-109, queries fan out and results are folded
// by the proxy (framework/proxy.hpp:249-266, aggregators.hpp:27-63), MIX is driven from
// framework/mixer/push_mixer.cpp:410-472 and linear_mixer.cpp:547-561,613-628.  This file puts
// the corresponding exchanges under the same C-ABI, so that a C++ host reaches them without any
// Python:
//   * a sharded query = shard-local scan -> ONE ncclAllGather of the packed (key, global index)
//     candidates -> merge, identical on every rank;
//   * a MIX round = one small all-gather of counts + ONE all-gather of the payload (signature
//     records packed on the device + the host's key / version blob), then a scatter of the rows
//     the host's version arbitration admits;
//   * two shapes of host: one process per GPU (jubann_comm_unique_id / jubann_comm_init, the
//     id travels by whatever the host has: the Jubatus server's ZooKeeper / RPC, torchrun's
//     store), and one process driving several GPUs (jubann_create_multi: a handle that shards
//     over a device list and answers the ordinary host-pointer entry points).
//
// NCCL is opened with dlopen on first use instead of being a link-time dependency: a process
// that has torch loaded already holds torch's own libnccl.so.2, and two different NCCL builds
// under one SONAME in one process is what this avoids (the first one loaded serves both).
// Every operand of every collective and of every bulk copy is this GPU's own memory; the
// cross-GPU traffic is NCCL's.
#include "handle.h"

#include <dlfcn.h>
#include <nccl.h>
#include <stdio.h>
#include <string.h>

#include <thread>

using namespace jbn;

namespace {

struct NcclApi {
  void* dl = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;
};

std::mutex g_nccl_mu;
NcclApi g_nccl;
bool g_nccl_tried = false;

const NcclApi* nccl_api() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl_tried) return g_nccl.dl ? &g_nccl : nullptr;
  g_nccl_tried = true;
  const char* names[] = {getenv("JUBANN_NCCL_LIBRARY"), "libnccl.so.2", "libnccl.so"};
  void* dl = nullptr;
  for (const char* n : names) {
    if (!n || !*n) continue;
    dl = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (dl) break;
  }
  if (!dl) {
    const char* e = dlerror();
    g_nccl.error = std::string("cannot load NCCL (libnccl.so.2): ") + (e ? e : "?");
    return nullptr;
  }
#define NCCL_SYM(field, name)                                                    \
  *reinterpret_cast<void**>(&g_nccl.field) = dlsym(dl, name);                    \
  if (!g_nccl.field) { g_nccl.error = std::string("NCCL symbol not found: ") + name; return nullptr; }
  NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
  NCCL_SYM(CommInitRank, "ncclCommInitRank");
  NCCL_SYM(CommDestroy, "ncclCommDestroy");
  NCCL_SYM(AllGather, "ncclAllGather");
  NCCL_SYM(Broadcast, "ncclBroadcast");
  NCCL_SYM(GroupStart, "ncclGroupStart");
  NCCL_SYM(GroupEnd, "ncclGroupEnd");
  NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef NCCL_SYM
  g_nccl.dl = dl;
  return &g_nccl;
}

int nccl_fail(jubann_t* h, const NcclApi* api, ncclResult_t r, const char* what) {
  // like a CUDA error: the communicator is not usable afterwards
  h->cuda_failed = true;
  h->sticky = std::string(what) + ": " + (api ? api->GetErrorString(r) : "NCCL unavailable");
  return fail(JUBANN_E_CUDA, "%s", h->sticky.c_str());
}

#define NK(call)                                                         \
  do {                                                                   \
    ncclResult_t r_ = (call);                                            \
    if (r_ != ncclSuccess) return nccl_fail(h, api, r_, #call);          \
  } while (0)

size_t up16(size_t x) { return (x + 15) & ~(size_t)15; }

// a rank's candidates for (nq, k), one contiguous block: [keys u32 nq*k | pad to 16 | global index u64 nq*k]
size_t packed_bytes(uint64_t nq, uint32_t k) { return up16(nq * k * 4) + nq * (size_t)k * 8; }

}  // namespace

namespace jbn {

struct CommState {
  ncclComm_t comm = nullptr;
  uint32_t rank = 0, world = 1;
  DevBuf send, recv;          // packed candidates: this rank's block, all ranks' blocks
  DevBuf qbuf;                // group handles: [query signatures | query norms] on this shard
  // MIX round state
  DevBuf mix_send, mix_recv, mix_counts, mix_sel;
  PinBuf mix_host;            // counts, then every rank's blob
  uint64_t mix_slot_bytes = 0, mix_rec_off = 0, mix_blob_off = 0;
  std::vector<uint64_t> mix_rows;
  uint64_t mix_bytes_last = 0;
};

void comm_destroy(jubann_t* h) {
  for (jubann_t* s : h->shards) jubann_destroy(s);
  h->shards.clear();
  CommState* c = h->comm;
  if (!c) return;
  if (c->comm) {   // (NCCL is looked up only by a handle that has a communicator: no need to load it to tear down a world of one)
    const NcclApi* api = nccl_api();
    if (api) api->CommDestroy(c->comm);
  }
  DevBuf* bufs[] = {&c->send, &c->recv, &c->qbuf, &c->mix_send, &c->mix_recv, &c->mix_counts, &c->mix_sel};
  for (DevBuf* b : bufs) b->release();
  c->mix_host.release();
  delete c;
  h->comm = nullptr;
}

static CommState* comm_state(jubann_t* h) {
  if (!h->comm) h->comm = new CommState;
  return h->comm;
}

// ---- one rank of a sharded table: scan -> all-gather -> merge, all on stream s ---------------
static int shard_scan_packed(jubann_t* h, uint64_t nq, const uint64_t* d_qsigs, const float* d_qnorms, uint32_t k,
                             cudaStream_t s) {
  CommState* c = comm_state(h);
  const size_t part = packed_bytes(nq, k);
  CK(c->send.ensure(part));
  CK(c->recv.ensure(part * c->world));
  unsigned char* base = c->send.as<unsigned char>();
  return query_device(h, nq, d_qsigs, d_qnorms, k, 0, nullptr, nullptr, nullptr, reinterpret_cast<uint32_t*>(base),
                      reinterpret_cast<uint64_t*>(base + up16(nq * k * 4)), s);
}

static int shard_allgather(jubann_t* h, const NcclApi* api, uint64_t nq, uint32_t k, cudaStream_t s) {
  CommState* c = h->comm;
  const size_t part = packed_bytes(nq, k);
  if (c->world == 1) {
    CK(cudaMemcpyAsync(c->recv.p, c->send.p, part, cudaMemcpyDeviceToDevice, s));
    return 0;
  }
  NK(api->AllGather(c->send.p, c->recv.p, part, ncclChar, c->comm, s));
  return 0;
}

static int shard_merge(jubann_t* h, uint64_t nq, uint32_t k, const float* d_qnorms, uint32_t flags,
                       uint64_t* d_out_index, float* d_out_score, uint32_t* d_out_count, cudaStream_t s) {
  CommState* c = h->comm;
  const size_t part = packed_bytes(nq, k);
  MergeOut o{};
  o.out_index = d_out_index; o.out_score = d_out_score; o.out_count = d_out_count;
  o.kout = k; o.col0 = 0;
  o.hash_num = h->hash_num; o.euclid = (h->method == EUCLID_LSH) ? 1 : 0; o.flags = flags; o.qnorms = d_qnorms;
  const unsigned char* base = c->recv.as<unsigned char>();
  CK(launch_merge_parts(reinterpret_cast<const uint32_t*>(base), reinterpret_cast<const uint64_t*>(base + up16(nq * k * 4)),
                        (uint32_t)nq, k, c->world, o, s, part / 4, part / 8));
  h->kernel_launches += 1;
  return 0;
}

// The query of a handle that is one shard of a table whose shards share a communicator: every
// rank makes the same call with the same queries and gets the same answer.
int sharded_query_device(jubann_t* h, uint64_t nq, const uint64_t* d_qsigs, const float* d_qnorms, uint32_t k,
                         uint32_t flags, uint64_t* d_out_index, float* d_out_score, uint32_t* d_out_count,
                         cudaStream_t s) {
  if (nq == 0) return 0;
  CommState* c = h->comm;
  if (!c || !c->comm && c->world > 1)
    return fail(JUBANN_E_INVALID, "sharded handle without a communicator: call jubann_comm_init first (or answer through scan_local_dev + merge_dev)");
  if (c->world != h->shard_count || c->rank != h->shard_rank)
    return fail(JUBANN_E_INVALID, "communicator rank %u of %u does not match shard %u of %u", c->rank, c->world, h->shard_rank, h->shard_count);
  if (nq > 0x7FFFFFFFull) return fail(JUBANN_E_INVALID, "too many queries in one call");
  if (k == 0) { CK(cudaMemsetAsync(d_out_count, 0, nq * 4, s)); return 0; }
  const NcclApi* api = c->world > 1 ? nccl_api() : nullptr;
  if (c->world > 1 && !api) return fail(JUBANN_E_CUDA, "%s", g_nccl.error.c_str());
  int rc = shard_scan_packed(h, nq, d_qsigs, d_qnorms, k, s);
  if (rc) return rc;
  rc = shard_allgather(h, api, nq, k, s);
  if (rc) return rc;
  return shard_merge(h, nq, k, d_qnorms, flags, d_out_index, d_out_score, d_out_count, s);
}

// ---- one process, several GPUs ---------------------------------------------------------------
// The group handle owns no table; shards[s] is an ordinary handle on devices[s] holding the rows
// g with g % ndev == s.  The group's mutex serialises calls; each shard is driven on its own
// stream from this one thread, NCCL calls of one step are bracketed by ncclGroupStart/End.
static uint32_t owner_of(uint64_t g, uint32_t G) { return (uint32_t)(g % G); }

int group_set_option(jubann_t* h, const char* name, int64_t value) {
  for (jubann_t* s : h->shards) {
    int rc = jubann_set_option(s, name, value);
    if (rc) return rc;
  }
  return 0;
}

int group_clear(jubann_t* h) {
  for (jubann_t* s : h->shards) {
    int rc = jubann_clear(s);
    if (rc) return rc;
  }
  h->global_rows = 0;
  return 0;
}

int group_size(jubann_t* h, uint64_t* global_rows, uint64_t* local_rows) {
  if (global_rows) *global_rows = h->global_rows;
  if (local_rows) *local_rows = h->global_rows;   // all of them live in this process
  return 0;
}

// splits a host batch by owning shard and hands every shard its own rows
template <typename PerShard>
static int for_each_shard_subset(jubann_t* h, uint64_t n, const uint64_t* row_index, PerShard fn) {
  const uint32_t G = (uint32_t)h->shards.size();
  std::vector<std::vector<uint64_t> > pos(G);
  for (uint64_t r = 0; r < n; ++r)
    if (row_index[r] != IDX_SENTINEL) pos[owner_of(row_index[r], G)].push_back(r);
  // large batches: one host thread per shard (staging copies, hashing and scatter run side by side)
  uint64_t busy = 0;
  for (uint32_t s = 0; s < G; ++s) busy += pos[s].empty() ? 0 : 1;
  if (busy > 1 && n >= 4096) {
    std::vector<int> rcs(G, 0);
    std::vector<std::string> msgs(G);
    std::vector<std::thread> th;
    for (uint32_t s = 0; s < G; ++s) {
      if (pos[s].empty()) continue;
      th.emplace_back([&, s]() {
        rcs[s] = fn(s, pos[s]);
        if (rcs[s]) msgs[s] = jubann_last_error();
      });
    }
    for (std::thread& t : th) t.join();
    for (uint32_t s = 0; s < G; ++s)
      if (rcs[s]) return fail(rcs[s], "%s", msgs[s].c_str());
    return 0;
  }
  for (uint32_t s = 0; s < G; ++s) {
    if (pos[s].empty()) continue;
    int rc = fn(s, pos[s]);
    if (rc) return rc;
  }
  return 0;
}

static int group_check_indices(jubann_t* h, uint64_t n, const uint64_t* row_index, uint64_t* new_global) {
  uint64_t g = h->global_rows;
  for (uint64_t r = 0; r < n; ++r) {
    if (row_index[r] > g) return fail(JUBANN_E_RANGE, "row index %llu skips past the table end %llu", (unsigned long long)row_index[r], (unsigned long long)g);
    if (row_index[r] == g) ++g;
  }
  *new_global = g;
  return 0;
}

// after a successful batch every shard knows the table's new end (its local count follows)
static void group_commit(jubann_t* h, uint64_t new_global) {
  for (jubann_t* s : h->shards) {
    std::lock_guard<std::mutex> lk(s->mu);
    commit_global(s, new_global);
  }
  if (new_global > h->global_rows) h->global_rows = new_global;
}

int group_set_rows_csr(jubann_t* h, uint64_t n, const uint64_t* row_index, const uint64_t* indptr,
                       const uint32_t* seed32, const uint64_t* keyhash64, const float* vals) {
  uint64_t new_global;
  int rc = group_check_indices(h, n, row_index, &new_global);
  if (rc) return rc;
  std::vector<uint64_t> tmp(row_index, row_index + n);
  strike_repeated_rows(tmp.data(), n);
  rc = for_each_shard_subset(h, n, tmp.data(), [&](uint32_t s, const std::vector<uint64_t>& pos) -> int {
    // the shard's own CSR batch, in batch order
    std::vector<uint64_t> ri(pos.size()), ip(pos.size() + 1, 0);
    uint64_t nnz = 0;
    for (size_t j = 0; j < pos.size(); ++j) nnz += indptr[pos[j] + 1] - indptr[pos[j]];
    std::vector<uint32_t> sd(seed32 ? nnz : 0);
    std::vector<uint64_t> kh(keyhash64 ? nnz : 0);
    std::vector<float> vv(nnz);
    uint64_t at = 0;
    for (size_t j = 0; j < pos.size(); ++j) {
      const uint64_t b = indptr[pos[j]], e = indptr[pos[j] + 1];
      ri[j] = tmp[pos[j]];
      if (seed32) memcpy(sd.data() + at, seed32 + b, (e - b) * 4);
      if (keyhash64) memcpy(kh.data() + at, keyhash64 + b, (e - b) * 8);
      if (e > b) memcpy(vv.data() + at, vals + b, (e - b) * 4);
      at += e - b;
      ip[j + 1] = at;
    }
    return jubann_set_rows_csr(h->shards[s], pos.size(), ri.data(), ip.data(), seed32 ? sd.data() : nullptr,
                               keyhash64 ? kh.data() : nullptr, vv.data());
  });
  if (rc) return rc;
  group_commit(h, new_global);
  return 0;
}

int group_set_rows_sig(jubann_t* h, uint64_t n, const uint64_t* row_index, const uint64_t* sigs, const float* norms) {
  uint64_t new_global;
  int rc = group_check_indices(h, n, row_index, &new_global);
  if (rc) return rc;
  std::vector<uint64_t> tmp(row_index, row_index + n);
  strike_repeated_rows(tmp.data(), n);
  const uint32_t w64 = h->w64;
  rc = for_each_shard_subset(h, n, tmp.data(), [&](uint32_t s, const std::vector<uint64_t>& pos) -> int {
    std::vector<uint64_t> ri(pos.size()), sg(pos.size() * w64);
    std::vector<float> nm(norms ? pos.size() : 0);
    for (size_t j = 0; j < pos.size(); ++j) {
      ri[j] = tmp[pos[j]];
      memcpy(sg.data() + j * w64, sigs + pos[j] * w64, w64 * 8);
      if (norms) nm[j] = norms[pos[j]];
    }
    return jubann_set_rows_sig(h->shards[s], pos.size(), ri.data(), sg.data(), norms ? nm.data() : nullptr);
  });
  if (rc) return rc;
  group_commit(h, new_global);
  return 0;
}

int group_get_rows_sig(jubann_t* h, uint64_t n, const uint64_t* row_index, uint64_t* out_sigs, float* out_norms) {
  for (uint64_t r = 0; r < n; ++r)
    if (row_index[r] >= h->global_rows) return fail(JUBANN_E_RANGE, "row index %llu out of range", (unsigned long long)row_index[r]);
  const uint32_t w64 = h->w64;
  return for_each_shard_subset(h, n, row_index, [&](uint32_t s, const std::vector<uint64_t>& pos) -> int {
    std::vector<uint64_t> ri(pos.size()), sg(pos.size() * w64);
    std::vector<float> nm(pos.size());
    for (size_t j = 0; j < pos.size(); ++j) ri[j] = row_index[pos[j]];
    int rc = jubann_get_rows_sig(h->shards[s], pos.size(), ri.data(), sg.data(), nm.data());
    if (rc) return rc;
    for (size_t j = 0; j < pos.size(); ++j) {
      memcpy(out_sigs + pos[j] * w64, sg.data() + j * w64, w64 * 8);
      if (out_norms) out_norms[pos[j]] = nm[j];
    }
    return 0;
  });
}

int group_move_row(jubann_t* h, uint64_t src, uint64_t dst) {
  if (src >= h->global_rows || dst >= h->global_rows) return fail(JUBANN_E_RANGE, "row index out of range");
  if (src == dst) return 0;
  std::vector<uint64_t> sig(h->w64);
  float norm = 0.0f;
  int rc = group_get_rows_sig(h, 1, &src, sig.data(), &norm);
  if (rc) return rc;
  return group_set_rows_sig(h, 1, &dst, sig.data(), h->method == EUCLID_LSH ? &norm : nullptr);
}

int group_truncate(jubann_t* h, uint64_t new_global_rows) {
  if (new_global_rows > h->global_rows) return fail(JUBANN_E_RANGE, "truncate cannot grow the table");
  for (jubann_t* s : h->shards) {
    int rc = jubann_truncate(s, std::min<uint64_t>(new_global_rows, s->global_rows));
    if (rc) return rc;
  }
  h->global_rows = new_global_rows;
  return 0;
}

int group_hash_csr(jubann_t* h, uint64_t n, const uint64_t* indptr, const uint32_t* seed32, const uint64_t* keyhash64,
                   const float* vals, uint64_t* out_sigs, float* out_norms) {
  return jubann_hash_csr(h->shards[0], n, indptr, seed32, keyhash64, vals, out_sigs, out_norms);
}

// queries already on shard 0's device (qbuf: [sigs | norms]) -> every shard -> scan -> all-gather
// -> merge on shard 0 -> host
static int group_answer(jubann_t* h, uint64_t nq, uint32_t k, uint32_t flags, uint64_t* out_index, float* out_score,
                        uint32_t* out_count) {
  const uint32_t G = (uint32_t)h->shards.size();
  const NcclApi* api = G > 1 ? nccl_api() : nullptr;
  if (G > 1 && !api) return fail(JUBANN_E_CUDA, "%s", g_nccl.error.c_str());
  const bool euclid = h->method == EUCLID_LSH;
  const size_t sig_b = up16(nq * h->w64 * 8), qb = sig_b + up16(nq * 4);
  if (k == 0) { memset(out_count, 0, nq * 4); return 0; }
  int prev = 0;
  cudaGetDevice(&prev);
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  // 1. the queries reach every shard
  for (uint32_t s = 1; s < G; ++s) {
    jubann_t* sh = h->shards[s];
    cudaSetDevice(sh->device);
    CKH(sh, comm_state(sh)->qbuf.ensure(qb));
  }
  if (G > 1) {
    api->GroupStart();
    for (uint32_t s = 0; s < G; ++s) {
      jubann_t* sh = h->shards[s];
      cudaSetDevice(sh->device);
      ncclResult_t r = api->Broadcast(sh->comm->qbuf.p, sh->comm->qbuf.p, qb, ncclChar, 0, sh->comm->comm, sh->stream);
      if (r != ncclSuccess) { api->GroupEnd(); return nccl_fail(h, api, r, "ncclBroadcast(queries)"); }
    }
    ncclResult_t r = api->GroupEnd();
    if (r != ncclSuccess) return nccl_fail(h, api, r, "ncclGroupEnd(broadcast)");
  }
  // 2. every shard scans its rows
  for (uint32_t s = 0; s < G; ++s) {
    jubann_t* sh = h->shards[s];
    Guard g(sh);
    const unsigned char* q = sh->comm->qbuf.as<unsigned char>();
    int rc = shard_scan_packed(sh, nq, reinterpret_cast<const uint64_t*>(q), euclid ? reinterpret_cast<const float*>(q + sig_b) : nullptr,
                               k, sh->stream);
    if (rc) return rc;
  }
  // 3. one all-gather of the packed candidates
  if (G > 1) {
    const size_t part = packed_bytes(nq, k);
    api->GroupStart();
    for (uint32_t s = 0; s < G; ++s) {
      jubann_t* sh = h->shards[s];
      cudaSetDevice(sh->device);
      ncclResult_t r = api->AllGather(sh->comm->send.p, sh->comm->recv.p, part, ncclChar, sh->comm->comm, sh->stream);
      if (r != ncclSuccess) { api->GroupEnd(); return nccl_fail(h, api, r, "ncclAllGather(candidates)"); }
    }
    ncclResult_t r = api->GroupEnd();
    if (r != ncclSuccess) return nccl_fail(h, api, r, "ncclGroupEnd(all-gather)");
  } else {
    jubann_t* sh = h->shards[0];
    Guard g(sh);
    int rc = shard_allgather(sh, nullptr, nq, k, sh->stream);
    if (rc) return rc;
  }
  // 4. shard 0 merges and answers
  {
    jubann_t* sh = h->shards[0];
    Guard g(sh);
    CKH(sh, sh->b_out_idx.ensure(nq * (size_t)k * 8));
    CKH(sh, sh->b_out_score.ensure(nq * (size_t)k * 4));
    CKH(sh, sh->b_out_count.ensure(nq * 4));
    const unsigned char* q = sh->comm->qbuf.as<unsigned char>();
    int rc = shard_merge(sh, nq, k, euclid ? reinterpret_cast<const float*>(q + sig_b) : nullptr, flags, sh->b_out_idx.as<uint64_t>(),
                         sh->b_out_score.as<float>(), sh->b_out_count.as<uint32_t>(), sh->stream);
    if (rc) return rc;
    CKH(sh, cudaMemcpyAsync(out_index, sh->b_out_idx.p, nq * (size_t)k * 8, cudaMemcpyDeviceToHost, sh->stream));
    CKH(sh, cudaMemcpyAsync(out_score, sh->b_out_score.p, nq * (size_t)k * 4, cudaMemcpyDeviceToHost, sh->stream));
    CKH(sh, cudaMemcpyAsync(out_count, sh->b_out_count.p, nq * 4, cudaMemcpyDeviceToHost, sh->stream));
    CKH(sh, cudaStreamSynchronize(sh->stream));
  }
  // the other shards' streams are left running: their next use is stream-ordered
  return 0;
}

int group_query_sig(jubann_t* h, uint64_t nq, const uint64_t* qsigs, const float* qnorms, uint32_t k, uint32_t flags,
                    uint64_t* out_index, float* out_score, uint32_t* out_count) {
  const size_t sig_b = up16(nq * h->w64 * 8), qb = sig_b + up16(nq * 4);
  {
    jubann_t* sh = h->shards[0];
    Guard g(sh);
    CommState* c = comm_state(sh);
    CKH(sh, c->qbuf.ensure(qb));
    CKH(sh, cudaMemcpyAsync(c->qbuf.p, qsigs, nq * sh->w64 * 8, cudaMemcpyHostToDevice, sh->stream));
    if (qnorms && sh->method == EUCLID_LSH)
      CKH(sh, cudaMemcpyAsync(c->qbuf.as<unsigned char>() + sig_b, qnorms, nq * 4, cudaMemcpyHostToDevice, sh->stream));
  }
  return group_answer(h, nq, k, flags, out_index, out_score, out_count);
}

int group_query_csr(jubann_t* h, uint64_t nq, const uint64_t* indptr, const uint32_t* seed32, const uint64_t* keyhash64,
                    const float* vals, uint32_t k, uint32_t flags, uint64_t* out_index, float* out_score,
                    uint32_t* out_count) {
  const uint64_t CHUNK = 1u << 18;
  for (uint64_t r0 = 0; r0 < nq; r0 += CHUNK) {
    const uint64_t r1 = std::min(nq, r0 + CHUNK), m = r1 - r0;
    const size_t sig_b = up16(m * h->w64 * 8), qb = sig_b + up16(m * 4);
    {
      jubann_t* sh = h->shards[0];
      Guard g(sh);
      CommState* c = comm_state(sh);
      CKH(sh, c->qbuf.ensure(qb));
      CsrDev d;
      int rc = stage_csr(sh, r0, r1, indptr, seed32, keyhash64, vals, &d);
      if (rc) return rc;
      unsigned char* q = c->qbuf.as<unsigned char>();
      rc = hash_device(sh, m, d.nnz, d.indptr, d.seed, d.keyhash, d.vals, reinterpret_cast<uint64_t*>(q), reinterpret_cast<float*>(q + sig_b));
      if (rc) return rc;
    }
    int rc = group_answer(h, m, k, flags, out_index + r0 * k, out_score + r0 * k, out_count + r0);
    if (rc) return rc;
  }
  return 0;
}

int group_query_rows(jubann_t* h, uint64_t nq, const uint64_t* row_index, uint32_t k, uint32_t flags, uint64_t* out_index,
                     float* out_score, uint32_t* out_count) {
  // the stored signatures live on whichever shard owns each row: fetch, then query
  std::vector<uint64_t> sigs(nq * h->w64);
  std::vector<float> norms(nq);
  int rc = group_get_rows_sig(h, nq, row_index, sigs.data(), norms.data());
  if (rc) return rc;
  return group_query_sig(h, nq, sigs.data(), norms.data(), k, flags, out_index, out_score, out_count);
}

int group_stats(jubann_t* h, uint64_t* kernel_launches, uint64_t* scan_launches, uint64_t* scan_passes, double* scan_ms) {
  uint64_t kl = 0, sl = 0, sp = 0;
  double ms = 0.0;
  for (jubann_t* s : h->shards) {
    uint64_t a = 0, b = 0, c = 0;
    double d = 0.0;
    int rc = jubann_stats(s, &a, &b, &c, &d);
    if (rc) return rc;
    kl += a; sl += b; sp += c; ms = std::max(ms, d);
  }
  if (kernel_launches) *kernel_launches = kl;
  if (scan_launches) *scan_launches = sl;
  if (scan_passes) *scan_passes = sp;
  if (scan_ms) *scan_ms = ms;
  return 0;
}

int group_stats_reset(jubann_t* h, int time_scan_kernels) {
  for (jubann_t* s : h->shards) {
    int rc = jubann_stats_reset(s, time_scan_kernels);
    if (rc) return rc;
  }
  return 0;
}

}  // namespace jbn

extern "C" {

int jubann_comm_unique_id(void* id) {
  if (!id) return fail(JUBANN_E_INVALID, "null argument");
  static_assert(JUBANN_COMM_ID_BYTES == NCCL_UNIQUE_ID_BYTES, "id size");
  const NcclApi* api = nccl_api();
  if (!api) return fail(JUBANN_E_CUDA, "%s", g_nccl.error.c_str());
  ncclUniqueId u;
  ncclResult_t r = api->GetUniqueId(&u);
  if (r != ncclSuccess) return fail(JUBANN_E_CUDA, "ncclGetUniqueId: %s", api->GetErrorString(r));
  memcpy(id, &u, NCCL_UNIQUE_ID_BYTES);
  return 0;
}

int jubann_comm_init(jubann_t* h, const void* id, uint32_t rank, uint32_t world) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (!h->shards.empty()) return fail(JUBANN_E_INVALID, "a multi-device handle has its communicator already");
  if (world < 1 || rank >= world) return fail(JUBANN_E_INVALID, "invalid rank %u of %u", rank, world);
  if (world > 1 && !id) return fail(JUBANN_E_INVALID, "null argument");
  Guard g(h);
  if (h->comm && h->comm->comm) return fail(JUBANN_E_INVALID, "the handle already has a communicator");
  CommState* c = comm_state(h);
  c->rank = rank; c->world = world;
  if (world == 1) return 0;
  const NcclApi* api = nccl_api();
  if (!api) return fail(JUBANN_E_CUDA, "%s", g_nccl.error.c_str());
  ncclUniqueId u;
  memcpy(&u, id, NCCL_UNIQUE_ID_BYTES);
  NK(api->CommInitRank(&c->comm, (int)world, u, (int)rank));
  return 0;
}

int jubann_comm_info(jubann_t* h, uint32_t* rank, uint32_t* world) {
  int rc = check_handle(h);
  if (rc) return rc;
  Guard g(h);
  if (rank) *rank = h->comm ? h->comm->rank : 0;
  if (world) *world = !h->shards.empty() ? (uint32_t)h->shards.size() : (h->comm ? h->comm->world : 1);
  return 0;
}

int jubann_create_multi(const char* method, uint32_t hash_num, const int* devices, int ndev, jubann_t** out) {
  if (!out) return fail(JUBANN_E_INVALID, "out is null");
  *out = nullptr;
  if (!devices || ndev < 1 || ndev > 64) return fail(JUBANN_E_INVALID, "device list is empty or too long");
  for (int i = 0; i < ndev; ++i)
    for (int j = 0; j < i; ++j)
      if (devices[i] == devices[j]) return fail(JUBANN_E_INVALID, "device %d is listed twice", devices[i]);
  jubann_t* top = new jubann;
  top->device = devices[0];
  for (int s = 0; s < ndev; ++s) {
    jubann_t* sh = nullptr;
    int rc = jubann_create(method, hash_num, devices[s], (uint32_t)s, (uint32_t)ndev, &sh);
    if (rc) { comm_destroy(top); delete top; return rc; }
    top->shards.push_back(sh);
    CommState* c = comm_state(sh);
    c->rank = (uint32_t)s; c->world = (uint32_t)ndev;
  }
  top->method = top->shards[0]->method; top->hash_num = hash_num;
  top->w64 = top->shards[0]->w64; top->w32 = top->shards[0]->w32;
  top->num_sms = top->shards[0]->num_sms;
  if (ndev > 1) {
    const NcclApi* api = nccl_api();
    if (!api) { comm_destroy(top); delete top; return fail(JUBANN_E_CUDA, "%s", g_nccl.error.c_str()); }
    ncclUniqueId u;
    ncclResult_t r = api->GetUniqueId(&u);
    int prev = 0;
    cudaGetDevice(&prev);
    if (r == ncclSuccess) {
      api->GroupStart();
      for (int s = 0; s < ndev && r == ncclSuccess; ++s) {
        cudaSetDevice(devices[s]);
        r = api->CommInitRank(&top->shards[s]->comm->comm, ndev, u, s);
      }
      ncclResult_t r2 = api->GroupEnd();
      if (r == ncclSuccess) r = r2;
    }
    cudaSetDevice(prev);
    if (r != ncclSuccess) {
      fail(JUBANN_E_CUDA, "NCCL communicator over %d devices: %s", ndev, api->GetErrorString(r));
      comm_destroy(top); delete top;
      return JUBANN_E_CUDA;
    }
  }
  *out = top;
  return 0;
}

// ---- MIX --------------------------------------------------------------------------------------
int jubann_mix_allgather(jubann_t* h, uint64_t n, const uint64_t* row_index, const void* blob, uint64_t blob_bytes,
                         uint64_t* out_rows, uint64_t* out_blob_bytes, const void** out_blobs) {
  int rc = check_handle(h);
  if (rc) return rc;
  if ((n && !row_index) || (blob_bytes && !blob) || !out_rows || !out_blob_bytes || !out_blobs) return fail(JUBANN_E_INVALID, "null argument");
  Guard g(h);
  if (!h->shards.empty() || h->shard_count != 1) return fail(JUBANN_E_INVALID, "MIX runs between replicas (single-shard handles)");
  CommState* c = h->comm;
  if (!c) return fail(JUBANN_E_INVALID, "no communicator: call jubann_comm_init first");
  const NcclApi* api = c->world > 1 ? nccl_api() : nullptr;
  if (c->world > 1 && !api) return fail(JUBANN_E_CUDA, "%s", g_nccl.error.c_str());
  for (uint64_t r = 0; r < n; ++r)
    if (row_index[r] >= h->global_rows) return fail(JUBANN_E_RANGE, "row index %llu out of range", (unsigned long long)row_index[r]);
  const uint32_t W = c->world, rec = 8 + 8 * h->w64 + 8;
  cudaStream_t s = h->stream;
  // 1. counts: (rows, blob bytes) of every rank
  CK(c->mix_counts.ensure((size_t)(W + 1) * 16));
  CK(c->mix_host.ensure((size_t)W * 16));
  uint64_t* hc = c->mix_host.as<uint64_t>();
  hc[0] = n; hc[1] = blob_bytes;
  uint64_t* d_mine = c->mix_counts.as<uint64_t>();
  uint64_t* d_all = d_mine + 2;
  CK(cudaMemcpyAsync(d_mine, hc, 16, cudaMemcpyHostToDevice, s));
  if (W > 1) NK(api->AllGather(d_mine, d_all, 16, ncclChar, c->comm, s));
  else CK(cudaMemcpyAsync(d_all, d_mine, 16, cudaMemcpyDeviceToDevice, s));
  CK(cudaMemcpyAsync(hc, d_all, (size_t)W * 16, cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  uint64_t rows_max = 0, blob_max = 0;
  c->mix_rows.assign(W, 0);
  std::vector<uint64_t> blob_b(W);
  for (uint32_t r = 0; r < W; ++r) {
    c->mix_rows[r] = hc[2 * r]; blob_b[r] = hc[2 * r + 1];
    rows_max = std::max(rows_max, hc[2 * r]); blob_max = std::max(blob_max, hc[2 * r + 1]);
    out_rows[r] = hc[2 * r]; out_blob_bytes[r] = hc[2 * r + 1];
  }
  // 2. payload: [signature records of my rows | my blob], one slot per rank
  c->mix_rec_off = 0;
  c->mix_blob_off = up16(rows_max * rec);
  c->mix_slot_bytes = c->mix_blob_off + up16(blob_max);
  c->mix_bytes_last = c->mix_slot_bytes;
  if (c->mix_slot_bytes == 0) {
    for (uint32_t r = 0; r < W; ++r) out_blobs[r] = nullptr;
    return 0;
  }
  CK(c->mix_send.ensure(c->mix_slot_bytes));
  CK(c->mix_recv.ensure(c->mix_slot_bytes * W));
  CK(c->mix_host.ensure(c->mix_slot_bytes * W + blob_bytes + 64));
  unsigned char* hb = c->mix_host.as<unsigned char>();
  if (n) {
    CK(h->b_rowidx.ensure(n * 8));
    CK(cudaMemcpyAsync(h->b_rowidx.p, row_index, n * 8, cudaMemcpyHostToDevice, s));
    CK(launch_diff_pack(view(h), n, h->b_rowidx.as<uint64_t>(), c->mix_send.as<unsigned char>() + c->mix_rec_off, rec, 0, 1, h->num_sms, s));
    h->kernel_launches += 1;
  }
  if (blob_bytes) {
    unsigned char* stage = hb + c->mix_slot_bytes * W;   // pinned copy of the caller's blob
    memcpy(stage, blob, blob_bytes);
    CK(cudaMemcpyAsync(c->mix_send.as<unsigned char>() + c->mix_blob_off, stage, blob_bytes, cudaMemcpyHostToDevice, s));
  }
  if (W > 1) NK(api->AllGather(c->mix_send.p, c->mix_recv.p, c->mix_slot_bytes, ncclChar, c->comm, s));
  else CK(cudaMemcpyAsync(c->mix_recv.p, c->mix_send.p, c->mix_slot_bytes, cudaMemcpyDeviceToDevice, s));
  // the host needs the blobs only; the records stay on the device for jubann_mix_apply
  for (uint32_t r = 0; r < W; ++r) {
    out_blobs[r] = nullptr;
    if (!blob_b[r]) continue;
    unsigned char* dst = hb + (size_t)r * c->mix_slot_bytes + c->mix_blob_off;
    CK(cudaMemcpyAsync(dst, c->mix_recv.as<unsigned char>() + (size_t)r * c->mix_slot_bytes + c->mix_blob_off, blob_b[r],
                       cudaMemcpyDeviceToHost, s));
    out_blobs[r] = dst;
  }
  CK(cudaStreamSynchronize(s));
  return 0;
}

int jubann_mix_apply(jubann_t* h, uint64_t m, const uint32_t* src_rank, const uint32_t* src_pos, const uint64_t* dst_row) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (m == 0) return 0;
  if (!src_rank || !src_pos || !dst_row) return fail(JUBANN_E_INVALID, "null argument");
  Guard g(h);
  CommState* c = h->comm;
  if (!c || c->mix_rows.empty() || !c->mix_recv.p) return fail(JUBANN_E_INVALID, "no gathered MIX round to apply: call jubann_mix_allgather first");
  for (uint64_t e = 0; e < m; ++e) {
    if (src_rank[e] >= c->world || src_pos[e] >= c->mix_rows[src_rank[e]])
      return fail(JUBANN_E_RANGE, "MIX entry %llu names record %u of rank %u, which was not gathered", (unsigned long long)e, src_pos[e], src_rank[e]);
  }
  uint64_t new_global;
  rc = check_row_indices(h, m, dst_row, &new_global);
  if (rc) return rc;
  rc = reserve_global(h, new_global);
  if (rc) return rc;
  std::vector<uint64_t> dst(dst_row, dst_row + m);
  strike_repeated_rows(dst.data(), m);   // two replicas shipped the same key: the later entry wins
  cudaStream_t s = h->stream;
  const size_t bytes = m * 16;
  CK(c->mix_sel.ensure(bytes));
  unsigned char* d = c->mix_sel.as<unsigned char>();
  CK(cudaMemcpyAsync(d, dst.data(), m * 8, cudaMemcpyHostToDevice, s));
  CK(cudaMemcpyAsync(d + m * 8, src_rank, m * 4, cudaMemcpyHostToDevice, s));
  CK(cudaMemcpyAsync(d + m * 12, src_pos, m * 4, cudaMemcpyHostToDevice, s));
  TableView t = view(h);
  t.local_rows = std::max<uint64_t>(t.local_rows, new_global);
  CK(launch_mix_apply(t, m, c->mix_recv.as<unsigned char>() + c->mix_rec_off, c->mix_slot_bytes, 8 + 8 * h->w64 + 8,
                      reinterpret_cast<const uint32_t*>(d + m * 8), reinterpret_cast<const uint32_t*>(d + m * 12),
                      reinterpret_cast<const uint64_t*>(d), h->num_sms, s));
  h->kernel_launches += 1;
  CK(cudaStreamSynchronize(s));
  commit_global(h, new_global);
  return 0;
}

}  // extern "C"
