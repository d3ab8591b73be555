// The handle behind jubann_t and the helpers the translation units of libjubann.so that work on
// it share (capi.cu: single-GPU entry points; comm.cu: the collectives under the C-ABI).
#ifndef JUBANN_HANDLE_H_
#define JUBANN_HANDLE_H_

#include "../../include/jubann.h"
#include "internal.h"

#include <algorithm>
#include <string>
#include <vector>

namespace jbn {

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = std::max(bytes, (size_t)4096);
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <typename T> T* as() { return static_cast<T*>(p); }
};

struct PinBuf {   // pinned host memory, grow-only
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    size_t want = std::max(bytes, (size_t)4096);
    cudaError_t e = cudaMallocHost(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
  template <typename T> T* as() { return static_cast<T*>(p); }
};

struct CommState;   // comm.cu: NCCL communicator of a handle that is one rank of a group

}  // namespace jbn

struct jubann {
  std::mutex mu;
  int device = 0;
  jbn::Method method = jbn::LSH;
  uint32_t hash_num = 0, w64 = 0, w32 = 0;
  uint32_t shard_rank = 0, shard_count = 1;
  int num_sms = 148;
  cudaStream_t stream = nullptr;
  // Orders work the caller enqueued on its own stream through a _dev mutator before anything the
  // handle's stream does afterwards (host-pointer entry points, slab growth).
  cudaEvent_t dev_write_event = nullptr;
  // table
  void* slab = nullptr;
  float* norms = nullptr;
  uint64_t cap_rows = 0;     // local capacity (multiple of 32)
  uint64_t local_rows = 0;
  uint64_t global_rows = 0;
  // projection cache
  jbn::ProjCache cache{};
  bool cache_ready = false;
  uint32_t cache_slots_opt = 1u << 20;
  uint64_t cache_used_ub = 0;   // upper bound of the slots handed out (exact whenever the counters are read back)
  double* d_coslut = nullptr;
  uint32_t* h_counters = nullptr;  // pinned
  unsigned char* h_stage = nullptr;  // pinned, SMALL_STAGE_BYTES: small requests go through one copy each way
  // options
  uint32_t qp_opt = 0;
  int variant_opt = 0;
  int popc_variant_opt = 3;   // 7 carry-save adders, 9 POPC per 512-bit pair: best measured on B200 (1: 5 adders, 11 POPC)
  int interleave_opt = 1;
  int score_mode = 0;         // euclid_lsh: 0 = cos and score in double (frozen choice), 1 = cosf and score in float
  // staging (grow-only)
  jbn::DevBuf b_indptr, b_seed, b_keyhash, b_vals, b_rowidx, b_sigs, b_norms;
  jbn::DevBuf b_qsigs, b_qnorms, b_out_idx, b_out_score, b_out_count, b_cand, b_floor, b_tau, b_hist, b_stage;
  // collectives (comm.cu)
  jbn::CommState* comm = nullptr;
  std::vector<jubann*> shards;   // non-empty: a multi-device handle (jubann_create_multi); the table lives in the shards
  // stats
  uint64_t kernel_launches = 0, scan_launches = 0, scan_passes = 0;
  bool time_scan = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t> > scan_events;
  double scan_ms = 0.0;
  // sticky error
  bool cuda_failed = false;
  std::string sticky;
};

namespace jbn {

int fail(int code, const char* fmt, ...);
int cuda_fail(jubann_t* h, cudaError_t e, const char* what);
int check_handle(jubann_t* h);
TableView view(jubann_t* h);
uint64_t local_count(uint64_t global_rows, uint32_t rank, uint32_t G);

#define CKH(hh, call)                                               \
  do {                                                              \
    cudaError_t e_ = (call);                                        \
    if (e_ != cudaSuccess) return jbn::cuda_fail(hh, e_, #call);    \
  } while (0)
#define CK(call) CKH(h, call)

struct Guard {
  jubann_t* h;
  std::unique_lock<std::mutex> lk;
  int prev_dev = -1;
  explicit Guard(jubann_t* hh) : h(hh), lk(hh->mu) {
    cudaGetDevice(&prev_dev);
    if (prev_dev != h->device) cudaSetDevice(h->device);
  }
  ~Guard() {
    if (prev_dev >= 0 && prev_dev != h->device) cudaSetDevice(prev_dev);
  }
};

// ---- shared by capi.cu and comm.cu (defined in capi.cu; all expect the Guard to be held) ----
struct CsrDev {
  uint64_t nnz;
  const uint64_t* indptr; const uint32_t* seed; const uint64_t* keyhash; const float* vals;
};
// rows [r0, r1) of a host CSR batch -> device staging (indptr rebased to 0), on the handle's stream
int stage_csr(jubann_t* h, uint64_t r0, uint64_t r1, const uint64_t* indptr, const uint32_t* seed32,
              const uint64_t* keyhash64, const float* vals, CsrDev* out);
// CSR on the device -> signatures (+ norms) on the device
int hash_device(jubann_t* h, uint64_t n, uint64_t nnz, const uint64_t* d_indptr, const uint32_t* d_seed,
                const uint64_t* d_keyhash, const float* d_vals, uint64_t* d_sigs, float* d_norms,
                cudaStream_t st = nullptr, bool use_handle_stream = true);
// scan + list reduction; exactly one of (final outputs) / (raw candidates) is given
int query_device(jubann_t* h, uint64_t nq64, const uint64_t* d_qsigs, const float* d_qnorms, uint32_t k,
                 uint32_t flags, uint64_t* d_out_index, float* d_out_score, uint32_t* d_out_count,
                 uint32_t* d_cand_key, uint64_t* d_cand_index, cudaStream_t s);
// validates row indices (append or overwrite, no holes), returns the new global size
int check_row_indices(jubann_t* h, uint64_t n, const uint64_t* row_index, uint64_t* new_global);
// slab capacity for `new_global` global rows (sizes are not changed), then the commit of the sizes
int reserve_global(jubann_t* h, uint64_t new_global);
void commit_global(jubann_t* h, uint64_t new_global);
// Strikes out (IDX_SENTINEL) every entry of a batch of destination rows that a later entry of the
// same batch overwrites: the scatter kernels write one thread per word with no order between
// entries, so a row written twice in one launch could end up torn.  `idx` is the staged copy.
void strike_repeated_rows(uint64_t* idx, uint64_t n);
// after work was enqueued on the caller's stream `st` that writes the table: the handle's own
// stream (host-pointer entry points, slab growth) runs after it
int order_after_caller_stream(jubann_t* h, cudaStream_t st);

// ---- comm.cu ----
void comm_destroy(jubann_t* h);
// the query of one shard of a table whose shards share a communicator (every rank calls it)
int sharded_query_device(jubann_t* h, uint64_t nq, const uint64_t* d_qsigs, const float* d_qnorms, uint32_t k,
                         uint32_t flags, uint64_t* d_out_index, float* d_out_score, uint32_t* d_out_count,
                         cudaStream_t s);
// multi-device handles (jubann_create_multi): the host-pointer entry points of capi.cu forward here
int group_set_option(jubann_t* h, const char* name, int64_t value);
int group_clear(jubann_t* h);
int group_size(jubann_t* h, uint64_t* global_rows, uint64_t* local_rows);
int group_hash_csr(jubann_t* h, uint64_t n, const uint64_t* indptr, const uint32_t* seed32, const uint64_t* keyhash64,
                   const float* vals, uint64_t* out_sigs, float* out_norms);
int group_set_rows_csr(jubann_t* h, uint64_t n, const uint64_t* row_index, const uint64_t* indptr,
                       const uint32_t* seed32, const uint64_t* keyhash64, const float* vals);
int group_set_rows_sig(jubann_t* h, uint64_t n, const uint64_t* row_index, const uint64_t* sigs, const float* norms);
int group_get_rows_sig(jubann_t* h, uint64_t n, const uint64_t* row_index, uint64_t* out_sigs, float* out_norms);
int group_move_row(jubann_t* h, uint64_t src, uint64_t dst);
int group_truncate(jubann_t* h, uint64_t new_global_rows);
int group_query_sig(jubann_t* h, uint64_t nq, const uint64_t* qsigs, const float* qnorms, uint32_t k, uint32_t flags,
                    uint64_t* out_index, float* out_score, uint32_t* out_count);
int group_query_csr(jubann_t* h, uint64_t nq, const uint64_t* indptr, const uint32_t* seed32, const uint64_t* keyhash64,
                    const float* vals, uint32_t k, uint32_t flags, uint64_t* out_index, float* out_score,
                    uint32_t* out_count);
int group_query_rows(jubann_t* h, uint64_t nq, const uint64_t* row_index, uint32_t k, uint32_t flags, uint64_t* out_index,
                     float* out_score, uint32_t* out_count);
int group_stats(jubann_t* h, uint64_t* kernel_launches, uint64_t* scan_launches, uint64_t* scan_passes, double* scan_ms);
int group_stats_reset(jubann_t* h, int time_scan_kernels);

}  // namespace jbn
#endif
