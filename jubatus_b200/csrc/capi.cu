// extern "C" surface of libjubann.so (declared in include/jubann.h): handle, device
// memory, batching and error plumbing around the kernels in scan.cu / signature.cu /
// table.cu.  No CPU fallback: every entry point needs a CUDA device.
#include "handle.h"

#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <unordered_set>

using namespace jbn;

thread_local std::string g_last_error;

namespace jbn {
int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}

int cuda_fail(jubann_t* h, cudaError_t e, const char* what) {
  h->cuda_failed = true;
  h->sticky = std::string(what) + ": " + cudaGetErrorString(e);
  return fail(JUBANN_E_CUDA, "%s", h->sticky.c_str());
}

int check_handle(jubann_t* h) {
  if (!h) return fail(JUBANN_E_INVALID, "null handle");
  if (h->cuda_failed) return fail(JUBANN_E_CUDA, "handle is in a failed state: %s", h->sticky.c_str());
  return 0;
}

TableView view(jubann_t* h) {
  TableView t;
  t.slab = h->slab; t.norms = h->norms; t.local_rows = h->local_rows;
  t.w32 = h->w32; t.w64 = h->w64; t.hash_num = h->hash_num;
  return t;
}

uint64_t local_count(uint64_t global_rows, uint32_t rank, uint32_t G) {
  return global_rows > rank ? (global_rows - rank + G - 1) / G : 0;
}
}  // namespace jbn

namespace jbn {

// Requests whose inputs and results fit this many bytes take the latency path: one pinned
// staging buffer, one host->device copy, one device->host copy, one synchronisation.
static const size_t SMALL_STAGE_BYTES = 1u << 20;

// grow the slab (and norm column) to hold `rows` local rows; contents are preserved
int ensure_rows(jubann_t* h, uint64_t rows) {
  if (rows <= h->cap_rows) return 0;
  uint64_t cap = std::max<uint64_t>(h->cap_rows * 2, 1024);
  while (cap < rows) cap *= 2;
  cap = (cap + 31) / 32 * 32;
  void* nslab = nullptr;
  float* nnorms = nullptr;
  cudaError_t e = cudaMalloc(&nslab, slab_bytes(h->w32, cap));
  if (e != cudaSuccess) {
    // fall back to the exact size before giving up
    cap = (rows + 31) / 32 * 32;
    cudaGetLastError();
    e = cudaMalloc(&nslab, slab_bytes(h->w32, cap));
    if (e != cudaSuccess) { cudaGetLastError(); return fail(JUBANN_E_NOMEM, "cudaMalloc(%llu bytes) for the signature slab failed", (unsigned long long)slab_bytes(h->w32, cap)); }
  }
  CK(cudaMemsetAsync(nslab, 0, slab_bytes(h->w32, cap), h->stream));
  if (h->slab) CK(cudaMemcpyAsync(nslab, h->slab, slab_bytes(h->w32, h->cap_rows), cudaMemcpyDeviceToDevice, h->stream));
  if (h->method == EUCLID_LSH) {
    e = cudaMalloc(reinterpret_cast<void**>(&nnorms), cap * sizeof(float));
    if (e != cudaSuccess) { cudaGetLastError(); cudaFree(nslab); return fail(JUBANN_E_NOMEM, "cudaMalloc for the norm column failed"); }
    CK(cudaMemsetAsync(nnorms, 0, cap * sizeof(float), h->stream));
    if (h->norms) CK(cudaMemcpyAsync(nnorms, h->norms, h->cap_rows * sizeof(float), cudaMemcpyDeviceToDevice, h->stream));
  }
  CK(cudaStreamSynchronize(h->stream));
  if (h->slab) cudaFree(h->slab);
  if (h->norms) cudaFree(h->norms);
  h->slab = nslab; h->norms = nnorms; h->cap_rows = cap;
  return 0;
}

int ensure_cache(jubann_t* h) {
  if (h->cache_ready || h->method == MINHASH) return 0;
  ProjCache& c = h->cache;
  c.slots = h->cache_slots_opt;
  c.hash_num = h->hash_num;
  uint32_t ts = 1024;
  while (ts < 2ull * c.slots) ts <<= 1;
  c.table_size = ts;
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&c.table), (size_t)c.slots * c.hash_num * sizeof(float));
  if (e != cudaSuccess) { cudaGetLastError(); return fail(JUBANN_E_NOMEM, "cudaMalloc for the projection cache (%u slots x %u floats) failed", c.slots, c.hash_num); }
  CK(cudaMalloc(reinterpret_cast<void**>(&c.keys), (size_t)ts * 8));
  CK(cudaMalloc(reinterpret_cast<void**>(&c.slot_of), (size_t)ts * 4));
  CK(cudaMalloc(reinterpret_cast<void**>(&c.counters), 16));
  CK(cudaMalloc(reinterpret_cast<void**>(&c.fill_list), (size_t)c.slots * sizeof(uint2)));
  CK(cudaMemsetAsync(c.keys, 0, (size_t)ts * 8, h->stream));
  CK(cudaMemsetAsync(c.slot_of, 0, (size_t)ts * 4, h->stream));
  CK(cudaMemsetAsync(c.counters, 0, 16, h->stream));
  h->cache_ready = true;
  return 0;
}

void free_cache(jubann_t* h) {
  ProjCache& c = h->cache;
  if (c.table) cudaFree(c.table);
  if (c.keys) cudaFree(c.keys);
  if (c.slot_of) cudaFree(c.slot_of);
  if (c.counters) cudaFree(c.counters);
  if (c.fill_list) cudaFree(c.fill_list);
  c = ProjCache{};
  h->cache_ready = false;
  h->cache_used_ub = 0;
}

// CSR already on the device -> signatures (+norms) on the device
int hash_device(jubann_t* h, uint64_t n, uint64_t nnz, const uint64_t* d_indptr, const uint32_t* d_seed,
                const uint64_t* d_keyhash, const float* d_vals, uint64_t* d_sigs, float* d_norms,
                cudaStream_t st, bool use_handle_stream) {
  if (use_handle_stream) st = h->stream;
  int launches = 0;
  if (h->method == MINHASH) {
    CK(launch_signature_minhash(h->hash_num, n, d_indptr, d_keyhash, d_vals, d_sigs, d_norms, h->w64, h->num_sms, st));
    ++launches;
  } else {
    int rc = ensure_cache(h);
    if (rc) return rc;
    CK(launch_cache_resolve(h->cache, d_seed, nnz, h->num_sms, st, &launches));
    CK(cudaMemcpyAsync(h->h_counters, h->cache.counters, 16, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    // a full cache is not an error: features without a row are drawn on the spot by the
    // signature kernel, as the reference draws every feature (h_counters[2] counts them)
    const uint32_t nfill = h->h_counters[1];
    h->cache_used_ub = h->h_counters[0];
    if (nfill) {
      CK(launch_cache_fill(h->cache, nfill, h->num_sms, st, &launches));
      CK(cudaMemsetAsync(h->cache.counters + 1, 0, 4, st));
    }
    CK(launch_signature_lsh(h->cache, h->method == EUCLID_LSH, n, d_indptr, d_seed, d_vals, d_sigs, d_norms,
                            h->w64, h->num_sms, h->cache_used_ub >= h->cache.slots, st));
    ++launches;
  }
  h->kernel_launches += launches;
  return 0;
}

// stage rows [r0, r1) of a host CSR batch on the device (indptr rebased to 0)
int stage_csr(jubann_t* h, uint64_t r0, uint64_t r1, const uint64_t* indptr, const uint32_t* seed32,
              const uint64_t* keyhash64, const float* vals, CsrDev* out) {
  const uint64_t n = r1 - r0, base = indptr[r0], nnz = indptr[r1] - base;
  if (h->method == MINHASH) { if (!keyhash64 && nnz) return fail(JUBANN_E_INVALID, "minhash needs keyhash64"); }
  else if (!seed32 && nnz) return fail(JUBANN_E_INVALID, "lsh/euclid_lsh need seed32");
  if (!vals && nnz) return fail(JUBANN_E_INVALID, "vals is null");
  std::vector<uint64_t> rebased(n + 1);
  for (uint64_t i = 0; i <= n; ++i) {
    if (indptr[r0 + i] < base || (i && indptr[r0 + i] < indptr[r0 + i - 1])) return fail(JUBANN_E_INVALID, "indptr is not non-decreasing");
    rebased[i] = indptr[r0 + i] - base;
  }
  CK(h->b_indptr.ensure((n + 1) * 8));
  CK(cudaMemcpyAsync(h->b_indptr.p, rebased.data(), (n + 1) * 8, cudaMemcpyHostToDevice, h->stream));
  CK(cudaStreamSynchronize(h->stream));  // `rebased` is a temporary
  CK(h->b_vals.ensure(std::max<uint64_t>(nnz, 1) * 4));
  if (nnz) CK(cudaMemcpyAsync(h->b_vals.p, vals + base, nnz * 4, cudaMemcpyHostToDevice, h->stream));
  if (h->method == MINHASH) {
    CK(h->b_keyhash.ensure(std::max<uint64_t>(nnz, 1) * 8));
    if (nnz) CK(cudaMemcpyAsync(h->b_keyhash.p, keyhash64 + base, nnz * 8, cudaMemcpyHostToDevice, h->stream));
  } else {
    CK(h->b_seed.ensure(std::max<uint64_t>(nnz, 1) * 4));
    if (nnz) CK(cudaMemcpyAsync(h->b_seed.p, seed32 + base, nnz * 4, cudaMemcpyHostToDevice, h->stream));
  }
  out->nnz = nnz;
  out->indptr = h->b_indptr.as<uint64_t>();
  out->seed = h->b_seed.as<uint32_t>();
  out->keyhash = h->b_keyhash.as<uint64_t>();
  out->vals = h->b_vals.as<float>();
  return 0;
}

const uint64_t CSR_CHUNK_ROWS = 1u << 18;

// validates row indices (append or overwrite, no holes) and returns the new global size.
// A shard of a sharded table may be handed every row (it keeps its own) or only the rows it owns:
// what must not skip is the shard's own slots.
int check_row_indices(jubann_t* h, uint64_t n, const uint64_t* row_index, uint64_t* new_global) {
  uint64_t g = h->global_rows;
  if (h->shard_count == 1) {
    for (uint64_t r = 0; r < n; ++r) {
      if (row_index[r] > g) return fail(JUBANN_E_RANGE, "row index %llu skips past the table end %llu", (unsigned long long)row_index[r], (unsigned long long)g);
      if (row_index[r] == g) ++g;
    }
    *new_global = g;
    return 0;
  }
  const uint32_t G = h->shard_count, rank = h->shard_rank;
  uint64_t loc = h->local_rows;
  for (uint64_t r = 0; r < n; ++r) {
    const uint64_t idx = row_index[r];
    if (idx == IDX_SENTINEL) return fail(JUBANN_E_RANGE, "row index out of range");
    if (idx % G == rank) {
      const uint64_t i = idx / G;
      if (i > loc) return fail(JUBANN_E_RANGE, "row index %llu skips past the end of shard %u (%llu local rows)", (unsigned long long)idx, rank, (unsigned long long)loc);
      if (i == loc) ++loc;
    }
    if (idx >= g) g = idx + 1;
  }
  if (local_count(g, rank, G) != loc)
    return fail(JUBANN_E_RANGE, "rows owned by shard %u are missing from the batch (table would end at %llu)", rank, (unsigned long long)g);
  *new_global = g;
  return 0;
}

// Capacity first, sizes last: a set_rows call that fails half way (allocation, CUDA error) leaves
// global_rows / local_rows as they were, so no row that was never stored becomes queryable.
int reserve_global(jubann_t* h, uint64_t new_global) {
  const uint64_t nl = local_count(new_global, h->shard_rank, h->shard_count);
  if (nl > 0xFFFFFFF0ull) return fail(JUBANN_E_RANGE, "more than 2^32-16 rows per shard are not supported");
  return ensure_rows(h, nl);
}
void commit_global(jubann_t* h, uint64_t new_global) {
  if (new_global <= h->global_rows) return;
  h->global_rows = new_global;
  h->local_rows = local_count(new_global, h->shard_rank, h->shard_count);
}
// view of the table as it will be once `new_global` rows are committed (for the scatter of a
// batch that appends: the kernel only needs the slab, but keep local_rows truthful)
static TableView view_grown(jubann_t* h, uint64_t new_global) {
  TableView t = view(h);
  t.local_rows = std::max<uint64_t>(t.local_rows, local_count(new_global, h->shard_rank, h->shard_count));
  return t;
}

void strike_repeated_rows(uint64_t* idx, uint64_t n) {
  if (n < 2) return;
  bool increasing = true;
  for (uint64_t r = 1; r < n && increasing; ++r) increasing = idx[r] > idx[r - 1];
  if (increasing) return;   // the common batch: appends / distinct rows in order
  std::unordered_set<uint64_t> seen;
  seen.reserve((size_t)n * 2);
  for (uint64_t r = n; r-- > 0;)
    if (idx[r] != IDX_SENTINEL && !seen.insert(idx[r]).second) idx[r] = IDX_SENTINEL;   // a later entry wins
}

int order_after_caller_stream(jubann_t* h, cudaStream_t st) {
  if (st == h->stream) return 0;
  CK(cudaEventRecord(h->dev_write_event, st));
  CK(cudaStreamWaitEvent(h->stream, h->dev_write_event, 0));
  return 0;
}

void record_scan_begin(jubann_t* h, cudaStream_t s, cudaEvent_t* a, cudaEvent_t* b) {
  *a = *b = nullptr;
  if (!h->time_scan) return;
  if (cudaEventCreate(a) != cudaSuccess || cudaEventCreate(b) != cudaSuccess) { *a = *b = nullptr; return; }
  cudaEventRecord(*a, s);
}
void record_scan_end(jubann_t* h, cudaStream_t s, cudaEvent_t a, cudaEvent_t b) {
  if (!a) return;
  cudaEventRecord(b, s);
  h->scan_events.push_back(std::make_pair(a, b));
}

// cos(theta_d), theta_d = (float)(d * M_PI / (float)hash_num), evaluated by the host libm so that the
// table holds the very values the CPU engine computes per row: in double (score_mode 0, the frozen
// choice) or with cosf (score_mode 1, the SURVEY A.9 alternate; stored widened, which is exact)
cudaError_t upload_coslut(jubann_t* h) {
  std::vector<double> lut(h->hash_num + 1);
  const float denom = static_cast<float>(h->hash_num);
  for (uint32_t i = 0; i <= h->hash_num; ++i) {
    const float theta = static_cast<float>(static_cast<uint64_t>(i) * M_PI / denom);
    lut[i] = h->score_mode ? static_cast<double>(cosf(theta)) : cos(static_cast<double>(theta));
  }
  return cudaMemcpy(h->d_coslut, lut.data(), lut.size() * 8, cudaMemcpyHostToDevice);
}

// The query path.  All pointers are device pointers; work is enqueued on `s`.
// Exactly one of (final outputs) / (raw candidate outputs) is given.
int query_device(jubann_t* h, uint64_t nq64, const uint64_t* d_qsigs, const float* d_qnorms, uint32_t k,
                 uint32_t flags, uint64_t* d_out_index, float* d_out_score, uint32_t* d_out_count,
                 uint32_t* d_cand_key, uint64_t* d_cand_index, cudaStream_t s) {
  if (nq64 == 0) return 0;
  if (nq64 > 0x7FFFFFFFull) return fail(JUBANN_E_INVALID, "too many queries in one call");
  const uint32_t nq = (uint32_t)nq64;
  const bool euclid = (h->method == EUCLID_LSH);
  if (euclid && !d_qnorms) return fail(JUBANN_E_INVALID, "euclid_lsh queries need norms");
  if (d_cand_key) {
    if (k) {
      CK(cudaMemsetAsync(d_cand_key, 0xFF, (size_t)nq * k * 4, s));
      CK(cudaMemsetAsync(d_cand_index, 0xFF, (size_t)nq * k * 8, s));
    }
  } else {
    CK(cudaMemsetAsync(d_out_count, 0, (size_t)nq * 4, s));
  }
  const uint64_t keff64 = std::min<uint64_t>(k, h->local_rows);
  if (keff64 == 0) return 0;
  const uint32_t keff = (uint32_t)keff64;
  const TableView t = view(h);
  const uint32_t kpage = std::min(keff, max_k_per_pass());
  const uint32_t pages = (keff + kpage - 1) / kpage;
  ScanConfig cfg;
  if (!plan_scan(t, euclid, nq, kpage, h->variant_opt, h->popc_variant_opt, h->qp_opt, h->interleave_opt, h->num_sms, &cfg))
    return fail(JUBANN_E_INVALID, "cannot plan a scan for k=%u", kpage);
  cfg.score_mode = (uint32_t)h->score_mode;
  // queries per launch bounded by candidate + shared-histogram scratch (<= 256 MiB)
  const size_t per_query = (size_t)cfg.nchunks * kpage * 8;
  const size_t hist_per_query = hist_scratch_bytes(1, h->hash_num, euclid);
  uint32_t nq_launch = (uint32_t)std::max<size_t>(cfg.qp, ((size_t)256 << 20) / (per_query + hist_per_query));
  nq_launch = std::max(cfg.qp, nq_launch / cfg.qp * cfg.qp);
  nq_launch = std::min(nq_launch, (nq + cfg.qp - 1) / cfg.qp * cfg.qp);
  if (nq_launch / cfg.qp > 65535u) nq_launch = 65535u * cfg.qp;
  CK(h->b_cand.ensure((size_t)nq_launch * per_query));
  if (pages > 1) CK(h->b_floor.ensure((size_t)nq * 8));
  CK(h->b_tau.ensure((size_t)nq_launch * 8));
  CK(h->b_hist.ensure(hist_scratch_bytes(nq_launch, h->hash_num, euclid)));
  for (uint32_t q0 = 0, nql = 0; q0 < nq; q0 += nql) {
    nql = std::min(nq_launch, nq - q0);
    // keep every launch made of full groups (they run the register-resident kernels); a
    // remainder smaller than one group gets a launch of its own
    if (nql > cfg.qp && nql % cfg.qp) nql -= nql % cfg.qp;
    for (uint32_t p = 0; p < pages; ++p) {
      const uint32_t kthis = std::min(kpage, keff - p * kpage);
      const uint64_t* floor = (p == 0) ? nullptr : h->b_floor.as<uint64_t>() + q0;
      cudaEvent_t ea, eb;
      int scan_kernels = 0;   // sample pre-pass (when it runs) + scan
      record_scan_begin(h, s, &ea, &eb);
      CK(launch_scan(t, euclid, cfg, d_qsigs + (size_t)q0 * h->w64, d_qnorms ? d_qnorms + q0 : nullptr,
                     h->d_coslut, nql, kthis, floor, h->b_tau.as<uint64_t>(), h->b_hist.as<uint32_t>(), h->b_cand.as<uint64_t>(), s, &scan_kernels));
      record_scan_end(h, s, ea, eb);
      MergeOut o{};
      o.kout = k; o.col0 = p * kpage;
      o.shard_rank = h->shard_rank; o.shard_count = h->shard_count;
      o.hash_num = h->hash_num; o.euclid = euclid ? 1 : 0; o.flags = flags;
      o.qnorms = d_qnorms ? d_qnorms + q0 : nullptr;
      o.count_accumulate = (p > 0);
      o.last_key = (pages > 1) ? h->b_floor.as<uint64_t>() + q0 : nullptr;
      if (d_cand_key) {
        o.cand_key = d_cand_key + (size_t)q0 * k;
        o.cand_index = d_cand_index + (size_t)q0 * k;
      } else {
        o.out_index = d_out_index + (size_t)q0 * k;
        o.out_score = d_out_score + (size_t)q0 * k;
        o.out_count = d_out_count + q0;
      }
      int merge_kernels = 0;   // (the thresholds of the scan are spent: their scratch carries the reduction's flags)
      CK(launch_merge_local(h->b_cand.as<uint64_t>(), nql, cfg.nchunks, kthis, o, s, h->b_tau.as<uint32_t>(), &merge_kernels));
      h->kernel_launches += scan_kernels + merge_kernels;
      h->scan_launches += 1;
      h->scan_passes += (nql + cfg.qp - 1) / cfg.qp;
    }
  }
  return 0;
}

}  // namespace jbn

// multi-device handles forward to comm.cu under the group's own lock
#define GROUP_DISPATCH(expr)                              \
  if (!h->shards.empty()) {                               \
    std::lock_guard<std::mutex> group_lock(h->mu);        \
    return (expr);                                        \
  }
#define NOT_FOR_GROUPS()                                  \
  if (!h->shards.empty())                                 \
    return fail(JUBANN_E_INVALID, "device-pointer entry points address one GPU: use the shard handles or the host-pointer entry points")

// final answers for a query batch already on the device: the plain scan for a whole table, the
// collective one for a shard whose handle has a communicator
static int answer_device(jubann_t* h, uint64_t nq, const uint64_t* d_qsigs, const float* d_qnorms, uint32_t k,
                         uint32_t flags, uint64_t* d_out_index, float* d_out_score, uint32_t* d_out_count,
                         cudaStream_t s) {
  if (h->shard_count == 1)
    return query_device(h, nq, d_qsigs, d_qnorms, k, flags, d_out_index, d_out_score, d_out_count, nullptr, nullptr, s);
  return sharded_query_device(h, nq, d_qsigs, d_qnorms, k, flags, d_out_index, d_out_score, d_out_count, s);
}

namespace jbn {
int set_error(int code, const char* message) { return fail(code, "%s", message); }
int handle_info(jubann* h, HandleInfo* out) {
  int rc = check_handle(h);
  if (rc) return rc;
  NOT_FOR_GROUPS();
  out->table = view(h);
  out->device = h->device;
  out->euclid = (h->method == EUCLID_LSH);
  out->shard_rank = h->shard_rank; out->shard_count = h->shard_count;
  out->variant = h->variant_opt; out->popc_variant = h->popc_variant_opt; out->interleave = h->interleave_opt;
  out->qp = h->qp_opt; out->num_sms = h->num_sms; out->coslut = h->d_coslut; out->mu = &h->mu;
  out->score_mode = h->score_mode;
  return 0;
}
void handle_add_launches(jubann* h, uint64_t kernels, uint64_t scans, uint64_t passes) {
  h->kernel_launches += kernels; h->scan_launches += scans; h->scan_passes += passes;
}
}  // namespace jbn

extern "C" {

const char* version(void) { return "jubann-b200 0.1.0 (jubatus 0.9.2 nearest_neighbor backend, sm_100a)"; }
int jubann_abi_version(void) { return JUBANN_ABI_VERSION; }
const char* jubann_last_error(void) { return g_last_error.c_str(); }

int jubann_create(const char* method, uint32_t hash_num, int device, uint32_t shard_rank,
                  uint32_t shard_count, jubann_t** out) {
  if (!out) return fail(JUBANN_E_INVALID, "out is null");
  *out = nullptr;
  if (!method) return fail(JUBANN_E_INVALID, "method is null");
  Method m;
  if (!strcmp(method, "lsh")) m = LSH;
  else if (!strcmp(method, "euclid_lsh")) m = EUCLID_LSH;
  else if (!strcmp(method, "minhash")) m = MINHASH;
  else return fail(JUBANN_E_INVALID, "unsupported_method: %s", method);
  if (hash_num < 1) return fail(JUBANN_E_INVALID, "invalid_parameter: hash_num should be >= 1");
  if (hash_num > 1024) return fail(JUBANN_E_INVALID, "invalid_parameter: hash_num > 1024 is not supported by this build");
  if (shard_count < 1 || shard_rank >= shard_count) return fail(JUBANN_E_INVALID, "invalid shard %u of %u", shard_rank, shard_count);
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return fail(JUBANN_E_CUDA, "no CUDA device: %s (this library has no CPU path)", e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  }
  if (device < 0 || device >= ndev) return fail(JUBANN_E_INVALID, "device %d out of range (%d devices)", device, ndev);
  jubann_t* h = new jubann;
  h->device = device; h->method = m; h->hash_num = hash_num;
  h->w64 = (hash_num + 63) / 64; h->w32 = padded_w32(hash_num);
  h->shard_rank = shard_rank; h->shard_count = shard_count;
  int prev = 0;
  cudaGetDevice(&prev);
  bool ok = true;
  ok = ok && cudaSetDevice(device) == cudaSuccess;
  cudaDeviceProp prop;
  ok = ok && cudaGetDeviceProperties(&prop, device) == cudaSuccess;
  if (ok) h->num_sms = prop.multiProcessorCount;
  ok = ok && cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&h->dev_write_event, cudaEventDisableTiming) == cudaSuccess;
  ok = ok && cudaMallocHost(reinterpret_cast<void**>(&h->h_counters), 16) == cudaSuccess;
  ok = ok && cudaMallocHost(reinterpret_cast<void**>(&h->h_stage), SMALL_STAGE_BYTES) == cudaSuccess;
  ok = ok && h->b_stage.ensure(SMALL_STAGE_BYTES) == cudaSuccess;
  if (ok && m == EUCLID_LSH) {
    ok = ok && cudaMalloc(reinterpret_cast<void**>(&h->d_coslut), ((size_t)hash_num + 1) * 8) == cudaSuccess;
    ok = ok && upload_coslut(h) == cudaSuccess;
  }
  if (!ok) {
    cudaError_t le = cudaGetLastError();
    fail(JUBANN_E_CUDA, "CUDA initialisation failed on device %d: %s", device, cudaGetErrorString(le));
    cudaSetDevice(prev);
    delete h;
    return JUBANN_E_CUDA;
  }
  cudaSetDevice(prev);
  *out = h;
  return 0;
}

int jubann_destroy(jubann_t* h) {
  if (!h) return 0;
  if (!h->shards.empty()) {   // multi-device handle: the shards own everything
    comm_destroy(h);
    delete h;
    return 0;
  }
  {
    Guard g(h);
    cudaStreamSynchronize(h->stream);
    comm_destroy(h);
    if (h->dev_write_event) cudaEventDestroy(h->dev_write_event);
    if (h->slab) cudaFree(h->slab);
    if (h->norms) cudaFree(h->norms);
    free_cache(h);
    if (h->d_coslut) cudaFree(h->d_coslut);
    if (h->h_counters) cudaFreeHost(h->h_counters);
    if (h->h_stage) cudaFreeHost(h->h_stage);
    DevBuf* bufs[] = {&h->b_indptr, &h->b_seed, &h->b_keyhash, &h->b_vals, &h->b_rowidx, &h->b_sigs, &h->b_norms,
                      &h->b_qsigs, &h->b_qnorms, &h->b_out_idx, &h->b_out_score, &h->b_out_count, &h->b_cand, &h->b_floor, &h->b_tau, &h->b_hist, &h->b_stage};
    for (DevBuf* b : bufs) b->release();
    for (auto& ev : h->scan_events) { cudaEventDestroy(ev.first); cudaEventDestroy(ev.second); }
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
  return 0;
}

int jubann_set_option(jubann_t* h, const char* name, int64_t value) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (!name) return fail(JUBANN_E_INVALID, "option name is null");
  GROUP_DISPATCH(group_set_option(h, name, value));
  Guard g(h);
  if (!strcmp(name, "cache_slots")) {
    if (value < 1 || value > (1ll << 28)) return fail(JUBANN_E_INVALID, "cache_slots out of range");
    if (h->cache_ready) return fail(JUBANN_E_INVALID, "cache_slots must be set before the first row or query (or after clear)");
    h->cache_slots_opt = (uint32_t)value;
  } else if (!strcmp(name, "reserve_rows")) {
    if (value < 0) return fail(JUBANN_E_INVALID, "reserve_rows out of range");
    const uint64_t nl = local_count((uint64_t)value, h->shard_rank, h->shard_count);
    if (nl > h->cap_rows) {
      // exact-size allocation: no doubling for a table whose final size is known
      const uint64_t keep = h->cap_rows;
      h->cap_rows = keep;
      uint64_t want = (nl + 31) / 32 * 32;
      void* nslab = nullptr; float* nnorms = nullptr;
      if (cudaMalloc(&nslab, slab_bytes(h->w32, want)) != cudaSuccess) { cudaGetLastError(); return fail(JUBANN_E_NOMEM, "cudaMalloc for %llu reserved rows failed", (unsigned long long)want); }
      CK(cudaMemsetAsync(nslab, 0, slab_bytes(h->w32, want), h->stream));
      if (h->slab) CK(cudaMemcpyAsync(nslab, h->slab, slab_bytes(h->w32, keep), cudaMemcpyDeviceToDevice, h->stream));
      if (h->method == EUCLID_LSH) {
        if (cudaMalloc(reinterpret_cast<void**>(&nnorms), want * sizeof(float)) != cudaSuccess) { cudaGetLastError(); cudaFree(nslab); return fail(JUBANN_E_NOMEM, "cudaMalloc for the norm column failed"); }
        CK(cudaMemsetAsync(nnorms, 0, want * sizeof(float), h->stream));
        if (h->norms) CK(cudaMemcpyAsync(nnorms, h->norms, keep * sizeof(float), cudaMemcpyDeviceToDevice, h->stream));
      }
      CK(cudaStreamSynchronize(h->stream));
      if (h->slab) cudaFree(h->slab);
      if (h->norms) cudaFree(h->norms);
      h->slab = nslab; h->norms = nnorms; h->cap_rows = want;
    }
  } else if (!strcmp(name, "gaussian_form")) {   // SURVEY A.9 alternate: 0 = cos (frozen choice), 1 = sin
    if (value < 0 || value > 1) return fail(JUBANN_E_INVALID, "gaussian_form out of range (0..1)");
    if (h->cache_used_ub || h->global_rows) return fail(JUBANN_E_INVALID, "gaussian_form must be set before the first row or query");
    h->cache.gauss_form = (uint32_t)value;
  } else if (!strcmp(name, "euclid_score")) {    // SURVEY A.9 alternate: 0 = double (frozen choice), 1 = cosf and float
    if (value < 0 || value > 1) return fail(JUBANN_E_INVALID, "euclid_score out of range (0..1)");
    h->score_mode = (int)value;
    if (h->method == EUCLID_LSH) {
      CK(cudaStreamSynchronize(h->stream));   // (no scan of this handle is reading the table while it changes)
      CK(upload_coslut(h));
    }
  } else if (!strcmp(name, "queries_per_pass")) {
    if (value < 0 || value > 64) return fail(JUBANN_E_INVALID, "queries_per_pass out of range (0..64)");
    h->qp_opt = (uint32_t)value;
  } else if (!strcmp(name, "scan_variant")) {
    if (value < 0 || value > 2) return fail(JUBANN_E_INVALID, "scan_variant out of range (0..2)");
    h->variant_opt = (int)value;
  } else if (!strcmp(name, "tma_geometry")) {   // tuning: (tiles_per_warp << 8) | stages, 0 = automatic
    set_tma_geometry_override((uint32_t)(value >> 8), (uint32_t)(value & 0xff));
  } else if (!strcmp(name, "scan_order")) {
    if (value < 0 || value > 1) return fail(JUBANN_E_INVALID, "scan_order out of range (0..1)");
    h->interleave_opt = (int)value;
  } else if (!strcmp(name, "popc_variant")) {
    if (value < 0 || value > 4) return fail(JUBANN_E_INVALID, "popc_variant out of range (0..4)");
    h->popc_variant_opt = (int)value;
  } else {
    return fail(JUBANN_E_INVALID, "unknown option: %s", name);
  }
  return 0;
}

int jubann_clear(jubann_t* h) {
  int rc = check_handle(h);
  if (rc) return rc;
  GROUP_DISPATCH(group_clear(h));
  Guard g(h);
  CK(cudaStreamSynchronize(h->stream));
  if (h->slab) CK(cudaMemsetAsync(h->slab, 0, slab_bytes(h->w32, h->cap_rows), h->stream));
  if (h->norms) CK(cudaMemsetAsync(h->norms, 0, h->cap_rows * sizeof(float), h->stream));
  CK(cudaStreamSynchronize(h->stream));
  h->local_rows = 0;
  h->global_rows = 0;
  if (h->cache_ready) {   // a cache that ran full is rebuilt, so that the next table's features get the rows
    CK(cudaMemcpy(h->h_counters, h->cache.counters, 16, cudaMemcpyDeviceToHost));
    if (h->h_counters[0] >= h->cache.slots) free_cache(h);
  }
  return 0;
}

int jubann_size(jubann_t* h, uint64_t* global_rows, uint64_t* local_rows) {
  int rc = check_handle(h);
  if (rc) return rc;
  GROUP_DISPATCH(group_size(h, global_rows, local_rows));
  Guard g(h);
  if (global_rows) *global_rows = h->global_rows;
  if (local_rows) *local_rows = h->local_rows;
  return 0;
}

int jubann_hash_csr(jubann_t* h, uint64_t n, const uint64_t* indptr, const uint32_t* seed32,
                    const uint64_t* keyhash64, const float* vals, uint64_t* out_sigs, float* out_norms) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (n == 0) return 0;
  if (!indptr || !out_sigs) return fail(JUBANN_E_INVALID, "null argument");
  GROUP_DISPATCH(group_hash_csr(h, n, indptr, seed32, keyhash64, vals, out_sigs, out_norms));
  Guard g(h);
  for (uint64_t r0 = 0; r0 < n; r0 += CSR_CHUNK_ROWS) {
    const uint64_t r1 = std::min(n, r0 + CSR_CHUNK_ROWS), m = r1 - r0;
    CsrDev d;
    rc = stage_csr(h, r0, r1, indptr, seed32, keyhash64, vals, &d);
    if (rc) return rc;
    CK(h->b_sigs.ensure(m * h->w64 * 8));
    CK(h->b_norms.ensure(m * 4));
    rc = hash_device(h, m, d.nnz, d.indptr, d.seed, d.keyhash, d.vals, h->b_sigs.as<uint64_t>(), h->b_norms.as<float>());
    if (rc) return rc;
    CK(cudaMemcpyAsync(out_sigs + r0 * h->w64, h->b_sigs.p, m * h->w64 * 8, cudaMemcpyDeviceToHost, h->stream));
    if (out_norms) CK(cudaMemcpyAsync(out_norms + r0, h->b_norms.p, m * 4, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
  }
  return 0;
}

// A batch may name the same destination row twice (the host's set_rows with a repeated id, a MIX
// round in which two replicas updated one key): column_table::add applied in order lets the last
// one win.  Returns row_index itself when no row repeats, else a copy in `tmp` with the earlier
// entries struck out.
static const uint64_t* last_wins(const uint64_t* row_index, uint64_t n, std::vector<uint64_t>& tmp) {
  bool increasing = true;
  for (uint64_t r = 1; r < n && increasing; ++r) increasing = row_index[r] > row_index[r - 1];
  if (increasing) return row_index;
  tmp.assign(row_index, row_index + n);
  strike_repeated_rows(tmp.data(), n);
  return tmp.data();
}

static int set_rows_csr_small(jubann_t* h, uint64_t n, const uint64_t* row_index, uint64_t new_global,
                              const uint64_t* indptr, const uint32_t* seed32, const uint64_t* keyhash64,
                              const float* vals, bool* taken);
int jubann_set_rows_csr(jubann_t* h, uint64_t n, const uint64_t* row_index, const uint64_t* indptr,
                        const uint32_t* seed32, const uint64_t* keyhash64, const float* vals) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (n == 0) return 0;
  if (!row_index || !indptr) return fail(JUBANN_E_INVALID, "null argument");
  GROUP_DISPATCH(group_set_rows_csr(h, n, row_index, indptr, seed32, keyhash64, vals));
  Guard g(h);
  uint64_t new_global;
  rc = check_row_indices(h, n, row_index, &new_global);
  if (rc) return rc;
  rc = reserve_global(h, new_global);   // capacity only: the sizes move once every row is stored
  if (rc) return rc;
  std::vector<uint64_t> tmp;
  row_index = last_wins(row_index, n, tmp);
  {
    bool taken = false;
    rc = set_rows_csr_small(h, n, row_index, new_global, indptr, seed32, keyhash64, vals, &taken);
    if (rc || taken) return rc;
  }
  for (uint64_t r0 = 0; r0 < n; r0 += CSR_CHUNK_ROWS) {
    const uint64_t r1 = std::min(n, r0 + CSR_CHUNK_ROWS), m = r1 - r0;
    CsrDev d;
    rc = stage_csr(h, r0, r1, indptr, seed32, keyhash64, vals, &d);
    if (rc) return rc;
    CK(h->b_sigs.ensure(m * h->w64 * 8));
    CK(h->b_norms.ensure(m * 4));
    CK(h->b_rowidx.ensure(m * 8));
    CK(cudaMemcpyAsync(h->b_rowidx.p, row_index + r0, m * 8, cudaMemcpyHostToDevice, h->stream));
    rc = hash_device(h, m, d.nnz, d.indptr, d.seed, d.keyhash, d.vals, h->b_sigs.as<uint64_t>(), h->b_norms.as<float>());
    if (rc) return rc;
    CK(launch_scatter_rows(view_grown(h, new_global), m, h->b_rowidx.as<uint64_t>(), 0, h->b_sigs.as<uint64_t>(),
                           h->b_norms.as<float>(), h->shard_rank, h->shard_count, h->num_sms, h->stream));
    h->kernel_launches += 1;
    CK(cudaStreamSynchronize(h->stream));
  }
  commit_global(h, new_global);
  return 0;
}

int jubann_set_rows_sig(jubann_t* h, uint64_t n, const uint64_t* row_index, const uint64_t* sigs,
                        const float* norms) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (n == 0) return 0;
  if (!row_index || !sigs) return fail(JUBANN_E_INVALID, "null argument");
  GROUP_DISPATCH(group_set_rows_sig(h, n, row_index, sigs, norms));
  Guard g(h);
  uint64_t new_global;
  rc = check_row_indices(h, n, row_index, &new_global);
  if (rc) return rc;
  rc = reserve_global(h, new_global);
  if (rc) return rc;
  std::vector<uint64_t> tmp;
  row_index = last_wins(row_index, n, tmp);
  CK(h->b_sigs.ensure(n * h->w64 * 8));
  CK(h->b_rowidx.ensure(n * 8));
  CK(cudaMemcpyAsync(h->b_sigs.p, sigs, n * h->w64 * 8, cudaMemcpyHostToDevice, h->stream));
  CK(cudaMemcpyAsync(h->b_rowidx.p, row_index, n * 8, cudaMemcpyHostToDevice, h->stream));
  const float* dn = nullptr;
  if (norms && h->method == EUCLID_LSH) {
    CK(h->b_norms.ensure(n * 4));
    CK(cudaMemcpyAsync(h->b_norms.p, norms, n * 4, cudaMemcpyHostToDevice, h->stream));
    dn = h->b_norms.as<float>();
  }
  CK(launch_scatter_rows(view_grown(h, new_global), n, h->b_rowidx.as<uint64_t>(), 0, h->b_sigs.as<uint64_t>(), dn,
                         h->shard_rank, h->shard_count, h->num_sms, h->stream));
  h->kernel_launches += 1;
  CK(cudaStreamSynchronize(h->stream));
  commit_global(h, new_global);
  return 0;
}

int jubann_set_rows_sig_dev(jubann_t* h, uint64_t n, uint64_t first_row_index, const uint64_t* d_sigs,
                            const float* d_norms, void* stream) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (n == 0) return 0;
  if (!d_sigs) return fail(JUBANN_E_INVALID, "null argument");
  NOT_FOR_GROUPS();
  Guard g(h);
  if (first_row_index > h->global_rows) return fail(JUBANN_E_RANGE, "row index skips past the table end");
  const uint64_t new_global = std::max(h->global_rows, first_row_index + n);
  rc = reserve_global(h, new_global);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CK(launch_scatter_rows(view_grown(h, new_global), n, nullptr, first_row_index, d_sigs, h->method == EUCLID_LSH ? d_norms : nullptr,
                         h->shard_rank, h->shard_count, h->num_sms, st));
  h->kernel_launches += 1;
  commit_global(h, new_global);
  return order_after_caller_stream(h, st);
}

int jubann_get_rows_sig(jubann_t* h, uint64_t n, const uint64_t* row_index, uint64_t* out_sigs,
                        float* out_norms) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (n == 0) return 0;
  if (!row_index || !out_sigs) return fail(JUBANN_E_INVALID, "null argument");
  GROUP_DISPATCH(group_get_rows_sig(h, n, row_index, out_sigs, out_norms));
  Guard g(h);
  for (uint64_t r = 0; r < n; ++r)
    if (row_index[r] >= h->global_rows) return fail(JUBANN_E_RANGE, "row index %llu out of range", (unsigned long long)row_index[r]);
  CK(h->b_sigs.ensure(n * h->w64 * 8));
  CK(h->b_norms.ensure(n * 4));
  CK(h->b_rowidx.ensure(n * 8));
  CK(cudaMemcpyAsync(h->b_rowidx.p, row_index, n * 8, cudaMemcpyHostToDevice, h->stream));
  CK(launch_gather_rows(view(h), n, h->b_rowidx.as<uint64_t>(), h->b_sigs.as<uint64_t>(), h->b_norms.as<float>(),
                        h->shard_rank, h->shard_count, h->num_sms, h->stream));
  h->kernel_launches += 1;
  CK(cudaMemcpyAsync(out_sigs, h->b_sigs.p, n * h->w64 * 8, cudaMemcpyDeviceToHost, h->stream));
  if (out_norms) CK(cudaMemcpyAsync(out_norms, h->b_norms.p, n * 4, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

int jubann_move_row(jubann_t* h, uint64_t src_index, uint64_t dst_index) {
  int rc = check_handle(h);
  if (rc) return rc;
  GROUP_DISPATCH(group_move_row(h, src_index, dst_index));
  Guard g(h);
  if (src_index >= h->global_rows || dst_index >= h->global_rows) return fail(JUBANN_E_RANGE, "row index out of range");
  if (h->shard_count != 1) return fail(JUBANN_E_INVALID, "move_row is the single-shard fast path; sharded hosts use get_rows_sig + set_rows_sig");
  if (src_index == dst_index) return 0;
  CK(h->b_sigs.ensure(h->w64 * 8));
  CK(h->b_norms.ensure(4));
  CK(h->b_rowidx.ensure(16));
  const uint64_t idx[2] = {src_index, dst_index};
  CK(cudaMemcpyAsync(h->b_rowidx.p, idx, 16, cudaMemcpyHostToDevice, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  CK(launch_gather_rows(view(h), 1, h->b_rowidx.as<uint64_t>(), h->b_sigs.as<uint64_t>(), h->b_norms.as<float>(), 0, 1, h->num_sms, h->stream));
  CK(launch_scatter_rows(view(h), 1, h->b_rowidx.as<uint64_t>() + 1, 0, h->b_sigs.as<uint64_t>(),
                         h->method == EUCLID_LSH ? h->b_norms.as<float>() : nullptr, 0, 1, h->num_sms, h->stream));
  h->kernel_launches += 2;
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

int jubann_truncate(jubann_t* h, uint64_t new_global_rows) {
  int rc = check_handle(h);
  if (rc) return rc;
  GROUP_DISPATCH(group_truncate(h, new_global_rows));
  Guard g(h);
  if (new_global_rows > h->global_rows) return fail(JUBANN_E_RANGE, "truncate cannot grow the table");
  h->global_rows = new_global_rows;
  h->local_rows = local_count(new_global_rows, h->shard_rank, h->shard_count);
  return 0;
}

static int query_host_outputs(jubann_t* h, uint64_t nq, uint32_t k, uint64_t* out_index, float* out_score,
                              uint32_t* out_count) {
  if (k) {
    CK(cudaMemcpyAsync(out_index, h->b_out_idx.p, nq * k * 8, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(out_score, h->b_out_score.p, nq * k * 4, cudaMemcpyDeviceToHost, h->stream));
  }
  CK(cudaMemcpyAsync(out_count, h->b_out_count.p, nq * 4, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return 0;
}

static int ensure_query_outputs(jubann_t* h, uint64_t nq, uint32_t k) {
  CK(h->b_out_idx.ensure(std::max<uint64_t>(nq * k, 1) * 8));
  CK(h->b_out_score.ensure(std::max<uint64_t>(nq * k, 1) * 4));
  CK(h->b_out_count.ensure(nq * 4));
  return 0;
}

// ---- latency path ------------------------------------------------------------------------------
// One RPC = one query is how the server calls this library (nearest_neighbor_impl.cpp:75-102), so
// the fixed cost of a call matters as much as the scan.  A request that fits SMALL_STAGE_BYTES is
// packed into the pinned staging buffer, copied in one piece, answered into the device staging
// buffer and copied back in one piece: two copies and one synchronisation per call instead of
// seven copies and three synchronisations.  Layout (256-byte aligned parts):
//   in : CSR [indptr | vals | seed32 or keyhash64]   or   [qsigs | qnorms]
//   dev: qsigs, qnorms (when hashed on the device)
//   out: [out_index | out_score | out_count]
static size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

struct SmallPlan {
  size_t in_bytes, sig_off, norm_off, out_off, idx_off, score_off, count_off, out_bytes, total;
};
static SmallPlan small_plan(const jubann_t* h, uint64_t nq, uint32_t k, size_t in_bytes, bool dev_sigs) {
  SmallPlan p;
  p.in_bytes = up256(in_bytes);
  size_t o = p.in_bytes;
  p.sig_off = o; if (dev_sigs) o += up256(nq * h->w64 * 8);
  p.norm_off = o; if (dev_sigs) o += up256(nq * 4);
  p.out_off = o;
  p.idx_off = o; o += up256(nq * (size_t)k * 8);
  p.score_off = o; o += up256(nq * (size_t)k * 4);
  p.count_off = o; o += up256(nq * 4);
  p.out_bytes = o - p.out_off;
  p.total = o;
  return p;
}

// run the query and bring the results back
static int small_finish(jubann_t* h, const SmallPlan& p, uint64_t nq, const uint64_t* d_qsigs, const float* d_qnorms,
                        uint32_t k, uint32_t flags, uint64_t* out_index, float* out_score,
                        uint32_t* out_count) {
  unsigned char* d = h->b_stage.as<unsigned char>();
  int rc = answer_device(h, nq, d_qsigs, d_qnorms, k, flags, reinterpret_cast<uint64_t*>(d + p.idx_off),
                         reinterpret_cast<float*>(d + p.score_off), reinterpret_cast<uint32_t*>(d + p.count_off), h->stream);
  if (rc) return rc;
  CK(cudaMemcpyAsync(h->h_stage + p.out_off, d + p.out_off, p.out_bytes, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  if (k) {
    memcpy(out_index, h->h_stage + p.idx_off, nq * (size_t)k * 8);
    memcpy(out_score, h->h_stage + p.score_off, nq * (size_t)k * 4);
  }
  memcpy(out_count, h->h_stage + p.count_off, nq * 4);
  return 0;
}

static int query_sig_small(jubann_t* h, uint64_t nq, const uint64_t* qsigs, const float* qnorms, uint32_t k,
                           uint32_t flags, uint64_t* out_index, float* out_score, uint32_t* out_count, bool* taken) {
  const size_t sig_b = up256(nq * h->w64 * 8), norm_b = h->method == EUCLID_LSH ? up256(nq * 4) : 0;
  const SmallPlan p = small_plan(h, nq, k, sig_b + norm_b, false);
  *taken = p.total <= SMALL_STAGE_BYTES;
  if (!*taken) return 0;
  memcpy(h->h_stage, qsigs, nq * h->w64 * 8);
  if (norm_b) memcpy(h->h_stage + sig_b, qnorms, nq * 4);
  unsigned char* d = h->b_stage.as<unsigned char>();
  CK(cudaMemcpyAsync(d, h->h_stage, sig_b + norm_b, cudaMemcpyHostToDevice, h->stream));
  return small_finish(h, p, nq, reinterpret_cast<const uint64_t*>(d), norm_b ? reinterpret_cast<const float*>(d + sig_b) : nullptr,
                      k, flags, out_index, out_score, out_count);
}

static int query_csr_small(jubann_t* h, uint64_t nq, const uint64_t* indptr, const uint32_t* seed32,
                           const uint64_t* keyhash64, const float* vals, uint32_t k, uint32_t flags,
                           uint64_t* out_index, float* out_score, uint32_t* out_count, bool* taken) {
  *taken = false;
  const uint64_t base = indptr[0], nnz = indptr[nq] - base;
  if (indptr[nq] < base || nnz > SMALL_STAGE_BYTES / 4) return 0;
  const bool mh = h->method == MINHASH;
  const size_t ip_b = up256((nq + 1) * 8), val_b = up256(nnz * 4), key_b = up256(nnz * (mh ? 8 : 4));
  const SmallPlan p = small_plan(h, nq, k, ip_b + val_b + key_b, true);
  if (p.total > SMALL_STAGE_BYTES) return 0;
  if (mh) { if (!keyhash64 && nnz) return fail(JUBANN_E_INVALID, "minhash needs keyhash64"); }
  else if (!seed32 && nnz) return fail(JUBANN_E_INVALID, "lsh/euclid_lsh need seed32");
  if (!vals && nnz) return fail(JUBANN_E_INVALID, "vals is null");
  *taken = true;
  uint64_t* ip = reinterpret_cast<uint64_t*>(h->h_stage);
  for (uint64_t i = 0; i <= nq; ++i) {
    if (indptr[i] < base || (i && indptr[i] < indptr[i - 1])) return fail(JUBANN_E_INVALID, "indptr is not non-decreasing");
    ip[i] = indptr[i] - base;
  }
  if (nnz) {
    memcpy(h->h_stage + ip_b, vals + base, nnz * 4);
    if (mh) memcpy(h->h_stage + ip_b + val_b, keyhash64 + base, nnz * 8);
    else memcpy(h->h_stage + ip_b + val_b, seed32 + base, nnz * 4);
  }
  unsigned char* d = h->b_stage.as<unsigned char>();
  CK(cudaMemcpyAsync(d, h->h_stage, ip_b + val_b + key_b, cudaMemcpyHostToDevice, h->stream));
  const uint64_t* d_ip = reinterpret_cast<const uint64_t*>(d);
  const float* d_vals = reinterpret_cast<const float*>(d + ip_b);
  uint64_t* d_sigs = reinterpret_cast<uint64_t*>(d + p.sig_off);
  float* d_norms = reinterpret_cast<float*>(d + p.norm_off);
  int launches = 0;
  if (mh) {
    CK(launch_signature_minhash(h->hash_num, nq, d_ip, reinterpret_cast<const uint64_t*>(d + ip_b + val_b), d_vals, d_sigs,
                                d_norms, h->w64, h->num_sms, h->stream));
    ++launches;
  } else {
    int rc = ensure_cache(h);
    if (rc) return rc;
    const uint32_t* d_seed = reinterpret_cast<const uint32_t*>(d + ip_b + val_b);
    CK(launch_cache_resolve(h->cache, d_seed, nnz, h->num_sms, h->stream, &launches));
    CK(launch_cache_fill_queued(h->cache, nnz, h->num_sms, h->stream, &launches));
    if (nnz) CK(cudaMemsetAsync(h->cache.counters + 1, 0, 4, h->stream));
    h->cache_used_ub += nnz;   // the count is not read back on this path: every feature may have been new
    CK(launch_signature_lsh(h->cache, h->method == EUCLID_LSH, nq, d_ip, d_seed, d_vals, d_sigs, d_norms, h->w64, h->num_sms,
                            h->cache_used_ub >= h->cache.slots, h->stream));
    ++launches;
  }
  h->kernel_launches += launches;
  return small_finish(h, p, nq, d_sigs, d_norms, k, flags, out_index, out_score, out_count);
}

// set_row for a few rows (one RPC): CSR and row indices in one copy, the cache fill off the
// device-side count, one synchronisation.
static int set_rows_csr_small(jubann_t* h, uint64_t n, const uint64_t* row_index, uint64_t new_global,
                              const uint64_t* indptr, const uint32_t* seed32, const uint64_t* keyhash64,
                              const float* vals, bool* taken) {
  *taken = false;
  const uint64_t base = indptr[0], nnz = indptr[n] - base;
  if (indptr[n] < base || nnz > SMALL_STAGE_BYTES / 4) return 0;
  const bool mh = h->method == MINHASH;
  const size_t ip_b = up256((n + 1) * 8), val_b = up256(nnz * 4), key_b = up256(nnz * (mh ? 8 : 4)), ri_b = up256(n * 8);
  const size_t in_b = ip_b + val_b + key_b + ri_b;
  const size_t sig_off = in_b, norm_off = sig_off + up256(n * h->w64 * 8), total = norm_off + up256(n * 4);
  if (total > SMALL_STAGE_BYTES) return 0;
  if (mh) { if (!keyhash64 && nnz) return fail(JUBANN_E_INVALID, "minhash needs keyhash64"); }
  else if (!seed32 && nnz) return fail(JUBANN_E_INVALID, "lsh/euclid_lsh need seed32");
  if (!vals && nnz) return fail(JUBANN_E_INVALID, "vals is null");
  *taken = true;
  uint64_t* ip = reinterpret_cast<uint64_t*>(h->h_stage);
  for (uint64_t i = 0; i <= n; ++i) {
    if (indptr[i] < base || (i && indptr[i] < indptr[i - 1])) return fail(JUBANN_E_INVALID, "indptr is not non-decreasing");
    ip[i] = indptr[i] - base;
  }
  if (nnz) {
    memcpy(h->h_stage + ip_b, vals + base, nnz * 4);
    if (mh) memcpy(h->h_stage + ip_b + val_b, keyhash64 + base, nnz * 8);
    else memcpy(h->h_stage + ip_b + val_b, seed32 + base, nnz * 4);
  }
  memcpy(h->h_stage + ip_b + val_b + key_b, row_index, n * 8);
  unsigned char* d = h->b_stage.as<unsigned char>();
  CK(cudaMemcpyAsync(d, h->h_stage, in_b, cudaMemcpyHostToDevice, h->stream));
  const uint64_t* d_ip = reinterpret_cast<const uint64_t*>(d);
  const float* d_vals = reinterpret_cast<const float*>(d + ip_b);
  const uint64_t* d_ri = reinterpret_cast<const uint64_t*>(d + ip_b + val_b + key_b);
  uint64_t* d_sigs = reinterpret_cast<uint64_t*>(d + sig_off);
  float* d_norms = reinterpret_cast<float*>(d + norm_off);
  int launches = 0;
  if (mh) {
    CK(launch_signature_minhash(h->hash_num, n, d_ip, reinterpret_cast<const uint64_t*>(d + ip_b + val_b), d_vals, d_sigs,
                                d_norms, h->w64, h->num_sms, h->stream));
    ++launches;
  } else {
    int rc = ensure_cache(h);
    if (rc) return rc;
    const uint32_t* d_seed = reinterpret_cast<const uint32_t*>(d + ip_b + val_b);
    CK(launch_cache_resolve(h->cache, d_seed, nnz, h->num_sms, h->stream, &launches));
    CK(launch_cache_fill_queued(h->cache, nnz, h->num_sms, h->stream, &launches));
    if (nnz) CK(cudaMemsetAsync(h->cache.counters + 1, 0, 4, h->stream));
    h->cache_used_ub += nnz;
    CK(launch_signature_lsh(h->cache, h->method == EUCLID_LSH, n, d_ip, d_seed, d_vals, d_sigs, d_norms, h->w64, h->num_sms,
                            h->cache_used_ub >= h->cache.slots, h->stream));
    ++launches;
  }
  CK(launch_scatter_rows(view_grown(h, new_global), n, d_ri, 0, d_sigs, d_norms, h->shard_rank, h->shard_count, h->num_sms, h->stream));
  h->kernel_launches += launches + 1;
  CK(cudaStreamSynchronize(h->stream));
  commit_global(h, new_global);
  return 0;
}

// set_row followed at once by neighbor_row(id) of that row: one copy in, one hash, one copy out.
int jubann_set_row_query_csr(jubann_t* h, uint64_t row_index, uint64_t nnz, const uint32_t* seed32, const uint64_t* keyhash64,
                             const float* vals, uint32_t k, uint32_t flags, uint64_t* out_index, float* out_score,
                             uint32_t* out_count) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (!out_count || (k && (!out_index || !out_score)) || (nnz && !vals)) return fail(JUBANN_E_INVALID, "null argument");
  if (!h->shards.empty() || h->shard_count != 1)
    return fail(JUBANN_E_INVALID, "set_row_query needs a single-shard handle (sharded hosts call set_rows_csr, then query_rows)");
  Guard g(h);
  const bool mh = h->method == MINHASH;
  if (mh ? (nnz && !keyhash64) : (nnz && !seed32)) return fail(JUBANN_E_INVALID, "missing feature hashes for this method");
  uint64_t new_global;
  rc = check_row_indices(h, 1, &row_index, &new_global);
  if (rc) return rc;
  rc = reserve_global(h, new_global);
  if (rc) return rc;
  const size_t ip_b = up256(16), val_b = up256(nnz * 4), key_b = up256(nnz * (mh ? 8 : 4)), ri_b = up256(8);
  const SmallPlan p = small_plan(h, 1, k, ip_b + val_b + key_b + ri_b, true);
  if (p.total > SMALL_STAGE_BYTES) return fail(JUBANN_E_INVALID, "set_row_query: the row or k is too large for the one-copy path");
  uint64_t* ip = reinterpret_cast<uint64_t*>(h->h_stage);
  ip[0] = 0; ip[1] = nnz;
  if (nnz) {
    memcpy(h->h_stage + ip_b, vals, nnz * 4);
    if (mh) memcpy(h->h_stage + ip_b + val_b, keyhash64, nnz * 8);
    else memcpy(h->h_stage + ip_b + val_b, seed32, nnz * 4);
  }
  memcpy(h->h_stage + ip_b + val_b + key_b, &row_index, 8);
  unsigned char* d = h->b_stage.as<unsigned char>();
  CK(cudaMemcpyAsync(d, h->h_stage, p.in_bytes, cudaMemcpyHostToDevice, h->stream));
  const uint64_t* d_ip = reinterpret_cast<const uint64_t*>(d);
  const float* d_vals = reinterpret_cast<const float*>(d + ip_b);
  const uint64_t* d_ri = reinterpret_cast<const uint64_t*>(d + ip_b + val_b + key_b);
  uint64_t* d_sigs = reinterpret_cast<uint64_t*>(d + p.sig_off);
  float* d_norms = reinterpret_cast<float*>(d + p.norm_off);
  int launches = 0;
  if (mh) {
    CK(launch_signature_minhash(h->hash_num, 1, d_ip, reinterpret_cast<const uint64_t*>(d + ip_b + val_b), d_vals, d_sigs,
                                d_norms, h->w64, h->num_sms, h->stream));
    ++launches;
  } else {
    rc = ensure_cache(h);
    if (rc) return rc;
    const uint32_t* d_seed = reinterpret_cast<const uint32_t*>(d + ip_b + val_b);
    CK(launch_cache_resolve(h->cache, d_seed, nnz, h->num_sms, h->stream, &launches));
    CK(launch_cache_fill_queued(h->cache, nnz, h->num_sms, h->stream, &launches));
    if (nnz) CK(cudaMemsetAsync(h->cache.counters + 1, 0, 4, h->stream));
    h->cache_used_ub += nnz;
    CK(launch_signature_lsh(h->cache, h->method == EUCLID_LSH, 1, d_ip, d_seed, d_vals, d_sigs, d_norms, h->w64, h->num_sms,
                            h->cache_used_ub >= h->cache.slots, h->stream));
    ++launches;
  }
  CK(launch_scatter_rows(view_grown(h, new_global), 1, d_ri, 0, d_sigs, d_norms, 0, 1, h->num_sms, h->stream));
  h->kernel_launches += launches + 1;
  commit_global(h, new_global);   // the scan that follows on the same stream ranks the table with the row in it
  return small_finish(h, p, 1, d_sigs, d_norms, k, flags, out_index, out_score, out_count);
}

int jubann_query_sig(jubann_t* h, uint64_t nq, const uint64_t* qsigs, const float* qnorms, uint32_t k,
                     uint32_t flags, uint64_t* out_index, float* out_score, uint32_t* out_count) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (nq == 0) return 0;
  if (!qsigs || !out_count || (k && (!out_index || !out_score))) return fail(JUBANN_E_INVALID, "null argument");
  if (h->method == EUCLID_LSH && !qnorms) return fail(JUBANN_E_INVALID, "euclid_lsh queries need norms");
  GROUP_DISPATCH(group_query_sig(h, nq, qsigs, qnorms, k, flags, out_index, out_score, out_count));
  Guard g(h);
  {
    bool taken = false;
    rc = query_sig_small(h, nq, qsigs, qnorms, k, flags, out_index, out_score, out_count, &taken);
    if (rc || taken) return rc;
  }
  CK(h->b_qsigs.ensure(nq * h->w64 * 8));
  CK(cudaMemcpyAsync(h->b_qsigs.p, qsigs, nq * h->w64 * 8, cudaMemcpyHostToDevice, h->stream));
  const float* dqn = nullptr;
  if (h->method == EUCLID_LSH) {
    CK(h->b_qnorms.ensure(nq * 4));
    CK(cudaMemcpyAsync(h->b_qnorms.p, qnorms, nq * 4, cudaMemcpyHostToDevice, h->stream));
    dqn = h->b_qnorms.as<float>();
  }
  rc = ensure_query_outputs(h, nq, k);
  if (rc) return rc;
  rc = answer_device(h, nq, h->b_qsigs.as<uint64_t>(), dqn, k, flags, h->b_out_idx.as<uint64_t>(),
                     h->b_out_score.as<float>(), h->b_out_count.as<uint32_t>(), h->stream);
  if (rc) return rc;
  return query_host_outputs(h, nq, k, out_index, out_score, out_count);
}

int jubann_query_csr(jubann_t* h, uint64_t nq, const uint64_t* indptr, const uint32_t* seed32,
                     const uint64_t* keyhash64, const float* vals, uint32_t k, uint32_t flags,
                     uint64_t* out_index, float* out_score, uint32_t* out_count) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (nq == 0) return 0;
  if (!indptr || !out_count || (k && (!out_index || !out_score))) return fail(JUBANN_E_INVALID, "null argument");
  GROUP_DISPATCH(group_query_csr(h, nq, indptr, seed32, keyhash64, vals, k, flags, out_index, out_score, out_count));
  Guard g(h);
  {
    bool taken = false;
    rc = query_csr_small(h, nq, indptr, seed32, keyhash64, vals, k, flags, out_index, out_score, out_count, &taken);
    if (rc || taken) return rc;
  }
  for (uint64_t r0 = 0; r0 < nq; r0 += CSR_CHUNK_ROWS) {
    const uint64_t r1 = std::min(nq, r0 + CSR_CHUNK_ROWS), m = r1 - r0;
    CsrDev d;
    rc = stage_csr(h, r0, r1, indptr, seed32, keyhash64, vals, &d);
    if (rc) return rc;
    CK(h->b_qsigs.ensure(m * h->w64 * 8));
    CK(h->b_qnorms.ensure(m * 4));
    rc = hash_device(h, m, d.nnz, d.indptr, d.seed, d.keyhash, d.vals, h->b_qsigs.as<uint64_t>(), h->b_qnorms.as<float>());
    if (rc) return rc;
    rc = ensure_query_outputs(h, m, k);
    if (rc) return rc;
    rc = answer_device(h, m, h->b_qsigs.as<uint64_t>(), h->b_qnorms.as<float>(), k, flags, h->b_out_idx.as<uint64_t>(),
                       h->b_out_score.as<float>(), h->b_out_count.as<uint32_t>(), h->stream);
    if (rc) return rc;
    rc = query_host_outputs(h, m, k, out_index + r0 * k, out_score + r0 * k, out_count + r0);
    if (rc) return rc;
  }
  return 0;
}

int jubann_query_rows(jubann_t* h, uint64_t nq, const uint64_t* row_index, uint32_t k, uint32_t flags,
                      uint64_t* out_index, float* out_score, uint32_t* out_count) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (nq == 0) return 0;
  if (!row_index || !out_count || (k && (!out_index || !out_score))) return fail(JUBANN_E_INVALID, "null argument");
  GROUP_DISPATCH(group_query_rows(h, nq, row_index, k, flags, out_index, out_score, out_count));
  Guard g(h);
  if (h->shard_count != 1) return fail(JUBANN_E_INVALID, "query_rows needs a single-shard handle");
  for (uint64_t r = 0; r < nq; ++r)
    if (row_index[r] >= h->global_rows) return fail(JUBANN_E_RANGE, "row index %llu out of range", (unsigned long long)row_index[r]);
  {
    const SmallPlan p = small_plan(h, nq, k, nq * 8, true);
    if (p.total <= SMALL_STAGE_BYTES) {   // latency path: one copy each way
      memcpy(h->h_stage, row_index, nq * 8);
      unsigned char* d = h->b_stage.as<unsigned char>();
      CK(cudaMemcpyAsync(d, h->h_stage, nq * 8, cudaMemcpyHostToDevice, h->stream));
      uint64_t* d_sigs = reinterpret_cast<uint64_t*>(d + p.sig_off);
      float* d_norms = reinterpret_cast<float*>(d + p.norm_off);
      CK(launch_gather_rows(view(h), nq, reinterpret_cast<const uint64_t*>(d), d_sigs, d_norms, 0, 1, h->num_sms, h->stream));
      h->kernel_launches += 1;
      return small_finish(h, p, nq, d_sigs, d_norms, k, flags, out_index, out_score, out_count);
    }
  }
  CK(h->b_rowidx.ensure(nq * 8));
  CK(h->b_qsigs.ensure(nq * h->w64 * 8));
  CK(h->b_qnorms.ensure(nq * 4));
  CK(cudaMemcpyAsync(h->b_rowidx.p, row_index, nq * 8, cudaMemcpyHostToDevice, h->stream));
  CK(launch_gather_rows(view(h), nq, h->b_rowidx.as<uint64_t>(), h->b_qsigs.as<uint64_t>(), h->b_qnorms.as<float>(), 0, 1, h->num_sms, h->stream));
  h->kernel_launches += 1;
  rc = ensure_query_outputs(h, nq, k);
  if (rc) return rc;
  rc = query_device(h, nq, h->b_qsigs.as<uint64_t>(), h->b_qnorms.as<float>(), k, flags, h->b_out_idx.as<uint64_t>(),
                    h->b_out_score.as<float>(), h->b_out_count.as<uint32_t>(), nullptr, nullptr, h->stream);
  if (rc) return rc;
  return query_host_outputs(h, nq, k, out_index, out_score, out_count);
}

int jubann_query_sig_dev(jubann_t* h, uint64_t nq, const uint64_t* d_qsigs, const float* d_qnorms, uint32_t k,
                         uint32_t flags, uint64_t* d_out_index, float* d_out_score, uint32_t* d_out_count,
                         void* stream) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (nq == 0) return 0;
  if (!d_qsigs || !d_out_count || (k && (!d_out_index || !d_out_score))) return fail(JUBANN_E_INVALID, "null argument");
  NOT_FOR_GROUPS();
  Guard g(h);
  return answer_device(h, nq, d_qsigs, d_qnorms, k, flags, d_out_index, d_out_score, d_out_count,
                       static_cast<cudaStream_t>(stream));
}

int jubann_hash_csr_dev(jubann_t* h, uint64_t n, uint64_t nnz, const uint64_t* d_indptr, const uint32_t* d_seed32,
                        const uint64_t* d_keyhash64, const float* d_vals, uint64_t* d_out_sigs, float* d_out_norms,
                        void* stream) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (n == 0) return 0;
  if (!d_indptr || !d_out_sigs || !d_out_norms || (nnz && !d_vals)) return fail(JUBANN_E_INVALID, "null argument");
  if (h->method == MINHASH ? (nnz && !d_keyhash64) : (nnz && !d_seed32)) return fail(JUBANN_E_INVALID, "missing feature hashes for this method");
  NOT_FOR_GROUPS();
  Guard g(h);
  return hash_device(h, n, nnz, d_indptr, d_seed32, d_keyhash64, d_vals, d_out_sigs, d_out_norms,
                     static_cast<cudaStream_t>(stream), false);
}

// rows first_row_index + i * row_stride (stride 1: a contiguous run; stride = shard_count: the rows of one shard only)
static int set_rows_csr_dev_impl(jubann_t* h, uint64_t n, uint64_t first_row_index, uint64_t row_stride, uint64_t nnz,
                                 const uint64_t* d_indptr, const uint32_t* d_seed32, const uint64_t* d_keyhash64,
                                 const float* d_vals, void* stream) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (n == 0) return 0;
  if (!d_indptr || (nnz && !d_vals)) return fail(JUBANN_E_INVALID, "null argument");
  if (h->method == MINHASH ? (nnz && !d_keyhash64) : (nnz && !d_seed32)) return fail(JUBANN_E_INVALID, "missing feature hashes for this method");
  NOT_FOR_GROUPS();
  Guard g(h);
  if (row_stride != 1) {
    if (row_stride != h->shard_count || first_row_index % h->shard_count != h->shard_rank)
      return fail(JUBANN_E_INVALID, "a strided batch must step by the shard count and start on a row of this shard");
    if (first_row_index / h->shard_count > h->local_rows) return fail(JUBANN_E_RANGE, "row index skips past the end of this shard");
  } else if (first_row_index > h->global_rows) {
    return fail(JUBANN_E_RANGE, "row index skips past the table end");
  }
  const uint64_t new_global = std::max(h->global_rows, first_row_index + (n - 1) * row_stride + 1);
  rc = reserve_global(h, new_global);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CK(h->b_sigs.ensure(n * h->w64 * 8));
  CK(h->b_norms.ensure(n * 4));
  rc = hash_device(h, n, nnz, d_indptr, d_seed32, d_keyhash64, d_vals, h->b_sigs.as<uint64_t>(), h->b_norms.as<float>(), st, false);
  if (rc) return rc;
  CK(launch_scatter_rows(view_grown(h, new_global), n, nullptr, first_row_index, h->b_sigs.as<uint64_t>(),
                         h->method == EUCLID_LSH ? h->b_norms.as<float>() : nullptr, h->shard_rank, h->shard_count, h->num_sms, st,
                         row_stride));
  h->kernel_launches += 1;
  commit_global(h, new_global);
  return order_after_caller_stream(h, st);
}

int jubann_set_rows_csr_dev(jubann_t* h, uint64_t n, uint64_t first_row_index, uint64_t nnz, const uint64_t* d_indptr,
                            const uint32_t* d_seed32, const uint64_t* d_keyhash64, const float* d_vals, void* stream) {
  return set_rows_csr_dev_impl(h, n, first_row_index, 1, nnz, d_indptr, d_seed32, d_keyhash64, d_vals, stream);
}

int jubann_set_rows_csr_strided_dev(jubann_t* h, uint64_t n, uint64_t first_row_index, uint64_t row_stride, uint64_t nnz,
                                    const uint64_t* d_indptr, const uint32_t* d_seed32, const uint64_t* d_keyhash64,
                                    const float* d_vals, void* stream) {
  if (row_stride == 0) return fail(JUBANN_E_INVALID, "row_stride is zero");
  return set_rows_csr_dev_impl(h, n, first_row_index, row_stride, nnz, d_indptr, d_seed32, d_keyhash64, d_vals, stream);
}

int jubann_scan_local_dev(jubann_t* h, uint64_t nq, const uint64_t* d_qsigs, const float* d_qnorms, uint32_t k,
                          uint32_t* d_cand_key, uint64_t* d_cand_index, void* stream) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (nq == 0 || k == 0) return 0;
  if (!d_qsigs || !d_cand_key || !d_cand_index) return fail(JUBANN_E_INVALID, "null argument");
  NOT_FOR_GROUPS();
  Guard g(h);
  return query_device(h, nq, d_qsigs, d_qnorms, k, 0, nullptr, nullptr, nullptr, d_cand_key, d_cand_index,
                      static_cast<cudaStream_t>(stream));
}

int jubann_merge_dev(jubann_t* h, uint64_t nq, uint32_t k, uint32_t parts, const uint32_t* d_cand_key,
                     const uint64_t* d_cand_index, const float* d_qnorms, uint32_t flags, uint64_t* d_out_index,
                     float* d_out_score, uint32_t* d_out_count, void* stream) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (nq == 0) return 0;
  if (!d_out_count) return fail(JUBANN_E_INVALID, "null argument");
  NOT_FOR_GROUPS();
  Guard g(h);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (k == 0 || parts == 0) { CK(cudaMemsetAsync(d_out_count, 0, nq * 4, s)); return 0; }
  if (!d_cand_key || !d_cand_index || !d_out_index || !d_out_score) return fail(JUBANN_E_INVALID, "null argument");
  if (h->method == EUCLID_LSH && !d_qnorms) return fail(JUBANN_E_INVALID, "euclid_lsh queries need norms");
  MergeOut o{};
  o.out_index = d_out_index; o.out_score = d_out_score; o.out_count = d_out_count;
  o.kout = k; o.col0 = 0;
  o.hash_num = h->hash_num; o.euclid = (h->method == EUCLID_LSH) ? 1 : 0; o.flags = flags; o.qnorms = d_qnorms;
  CK(launch_merge_parts(d_cand_key, d_cand_index, (uint32_t)nq, k, parts, o, s));
  h->kernel_launches += 1;
  return 0;
}

int jubann_diff_record_bytes(jubann_t* h, uint32_t* bytes) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (!bytes) return fail(JUBANN_E_INVALID, "null argument");
  *bytes = 8 + 8 * h->w64 + 8;
  return 0;
}

int jubann_diff_pack_dev(jubann_t* h, uint64_t n, const uint64_t* d_row_index, void* d_records, void* stream) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (n == 0) return 0;
  if (!d_row_index || !d_records) return fail(JUBANN_E_INVALID, "null argument");
  NOT_FOR_GROUPS();
  Guard g(h);
  CK(launch_diff_pack(view(h), n, d_row_index, d_records, 8 + 8 * h->w64 + 8, h->shard_rank, h->shard_count, h->num_sms,
                      static_cast<cudaStream_t>(stream)));
  h->kernel_launches += 1;
  return 0;
}

int jubann_diff_apply_dev(jubann_t* h, uint64_t n, const void* d_records, void* stream) {
  int rc = check_handle(h);
  if (rc) return rc;
  if (n == 0) return 0;
  if (!d_records) return fail(JUBANN_E_INVALID, "null argument");
  NOT_FOR_GROUPS();
  Guard g(h);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CK(launch_diff_apply(view(h), n, d_records, 8 + 8 * h->w64 + 8, h->shard_rank, h->shard_count, h->num_sms, st));
  h->kernel_launches += 1;
  return order_after_caller_stream(h, st);
}

int jubann_stats(jubann_t* h, uint64_t* kernel_launches, uint64_t* scan_launches, uint64_t* scan_passes,
                 double* scan_ms) {
  int rc = check_handle(h);
  if (rc) return rc;
  GROUP_DISPATCH(group_stats(h, kernel_launches, scan_launches, scan_passes, scan_ms));
  Guard g(h);
  for (auto& ev : h->scan_events) {
    float ms = 0.0f;
    cudaEventSynchronize(ev.second);
    if (cudaEventElapsedTime(&ms, ev.first, ev.second) == cudaSuccess) h->scan_ms += ms;
    cudaEventDestroy(ev.first);
    cudaEventDestroy(ev.second);
  }
  h->scan_events.clear();
  if (kernel_launches) *kernel_launches = h->kernel_launches;
  if (scan_launches) *scan_launches = h->scan_launches;
  if (scan_passes) *scan_passes = h->scan_passes;
  if (scan_ms) *scan_ms = h->scan_ms;
  return 0;
}

int jubann_stats_reset(jubann_t* h, int time_scan_kernels) {
  int rc = check_handle(h);
  if (rc) return rc;
  GROUP_DISPATCH(group_stats_reset(h, time_scan_kernels));
  Guard g(h);
  for (auto& ev : h->scan_events) { cudaEventDestroy(ev.first); cudaEventDestroy(ev.second); }
  h->scan_events.clear();
  h->kernel_launches = 0; h->scan_launches = 0; h->scan_passes = 0; h->scan_ms = 0.0;
  h->time_scan = time_scan_kernels != 0;
  return 0;
}

}  // extern "C"
