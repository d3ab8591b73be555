// jubann_query_sig_noalloc_dev: the query path for callers that must not allocate.
//
// Who this is for.  Some hosts cannot let a library call cudaMalloc once they are up:
// a server whose device memory is handed out once at start-up by its own pool; RPC
// threads that answer under a latency budget (cudaMalloc and cudaFree stall the whole
// device, so one thread's allocation delays every other thread's kernels); and processes
// in which some other thread may be recording work on a stream while queries arrive,
// where an unexpected allocation is not allowed to happen.  The ordinary query entry
// points grow a scratch buffer (the per-CTA candidate lists that the scan kernel writes and
// the merge kernel reads) with cudaMalloc the first time a larger batch shows up.  This one
// never does: the same scan and merge kernels run, but their scratch is the fixed array below.
//
// It has no counterpart in the reference (the CPU engine simply uses the heap).
//
// What you pay and what to watch for
//  * The array is part of the library image: every GPU that loads libjubann.so holds
//    JUBANN_NOALLOC_SCRATCH_BYTES (32 MiB) of device memory until the process exits,
//    whether or not this entry point is ever called.
//  * A call needs  nq * (chunks * min(k, rows) + 1) * 8 + nq * (hash_num + 33) * 4  bytes of it (second term: lsh / minhash only) (chunks = CTAs along the
//    table, at most 2 per SM).  A request that does not fit, or k above the single-pass
//    limit of 2048, is refused with JUBANN_E_INVALID and a message; nothing is launched.
//  * The first call on a device looks up the array's device address
//    (cudaGetSymbolAddress).  Make that first call during start-up, before any stream
//    capture is begun.
//  * All calls on a device share the one array, and it keeps whatever the last call left in
//    it.  The candidate slots need no clearing (every slot the merge kernel reads is written
//    by the scan kernel of the same call); the part that must start from a known state -- the
//    distance histogram and bound words the CTAs share -- is reset on the stream at the start
//    of every call (by the sample pre-pass kernel, or by cudaMemsetAsync when that kernel does
//    not run).  Two calls whose GPU work
//    overlaps (different streams, different handles or threads) would overwrite each other's
//    candidates -- run them on one stream or order them yourself.
//  * Inputs and outputs are the caller's device buffers, exactly as for
//    jubann_query_sig_dev; results are identical to that entry point's.
#include "../../include/jubann.h"
#include "internal.h"

#include <stdio.h>

#include <algorithm>

using namespace jbn;

static __device__ __align__(256) unsigned char g_noalloc_scratch[JUBANN_NOALLOC_SCRATCH_BYTES];

namespace {

const int kMaxDevices = 64;
std::mutex g_setup_mu;
void* g_scratch_ptr[kMaxDevices];   // device address of g_noalloc_scratch, per device ordinal
bool g_scratch_ready[kMaxDevices];

// Host code cannot use the name g_noalloc_scratch as a device pointer: here it is only a
// stand-in for the symbol.  Ask the runtime for the real address once per device, with that
// device current.
int scratch_address(int device, void** out) {
  if (device < 0 || device >= kMaxDevices) return set_error(JUBANN_E_INVALID, "device ordinal out of range for the no-alloc scratch");
  std::lock_guard<std::mutex> lk(g_setup_mu);
  if (!g_scratch_ready[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr) {
      char msg[256];
      snprintf(msg, sizeof(msg), "cudaGetSymbolAddress(no-alloc scratch) failed on device %d: %s", device,
               cudaGetErrorString(e));
      cudaGetLastError();
      return set_error(JUBANN_E_CUDA, msg);
    }
    g_scratch_ptr[device] = p;
    g_scratch_ready[device] = true;
  }
  *out = g_scratch_ptr[device];
  return 0;
}

struct DeviceScope {
  int prev = -1, want;
  explicit DeviceScope(int d) : want(d) {
    cudaGetDevice(&prev);
    if (prev != want) cudaSetDevice(want);
  }
  ~DeviceScope() {
    if (prev >= 0 && prev != want) cudaSetDevice(prev);
  }
};

int cuda_error(const char* what, cudaError_t e) {
  char msg[256];
  snprintf(msg, sizeof(msg), "%s: %s", what, cudaGetErrorString(e));
  return set_error(JUBANN_E_CUDA, msg);
}

}  // namespace

extern "C" int jubann_query_sig_noalloc_dev(jubann_t* h, uint64_t nq64, const uint64_t* d_qsigs, const float* d_qnorms,
                                            uint32_t k, uint32_t flags, uint64_t* d_out_index, float* d_out_score,
                                            uint32_t* d_out_count, void* stream) {
  HandleInfo hi;
  int rc = handle_info(h, &hi);
  if (rc) return rc;
  if (nq64 == 0) return 0;
  if (!d_qsigs || !d_out_count || (k && (!d_out_index || !d_out_score))) return set_error(JUBANN_E_INVALID, "null argument");
  if (nq64 > 0x7FFFFFFFull) return set_error(JUBANN_E_INVALID, "too many queries in one call");
  if (hi.shard_count != 1) return set_error(JUBANN_E_INVALID, "sharded handles answer through scan_local_dev + merge_dev");
  if (hi.euclid && !d_qnorms) return set_error(JUBANN_E_INVALID, "euclid_lsh queries need norms");
  const uint32_t nq = (uint32_t)nq64;
  std::lock_guard<std::mutex> lk(*hi.mu);
  rc = handle_info(h, &hi);   // re-read under the lock
  if (rc) return rc;
  DeviceScope dev(hi.device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);

  const uint64_t keff64 = std::min<uint64_t>(k, hi.table.local_rows);
  char msg[320];
  if (keff64 > max_k_per_pass()) {
    snprintf(msg, sizeof(msg), "no-alloc query: k=%u is above the single-pass limit of %u results", k, max_k_per_pass());
    return set_error(JUBANN_E_INVALID, msg);
  }
  ScanConfig cfg;
  const uint32_t keff = (uint32_t)keff64;
  if (keff) {
    if (!plan_scan(hi.table, hi.euclid, nq, keff, hi.variant, hi.popc_variant, hi.qp, hi.interleave, hi.num_sms, &cfg))
      return set_error(JUBANN_E_INVALID, "no-alloc query: cannot plan a scan for this k");
    cfg.score_mode = (uint32_t)hi.score_mode;
    const uint64_t need = (uint64_t)nq * cfg.nchunks * keff * 8ull + (uint64_t)nq * 8ull +   // lists + one threshold per query
                          (hi.euclid ? 0 : hist_scratch_bytes(nq, hi.table.hash_num, false));  // + shared distance histogram (hamming engines)
    if (need > (uint64_t)JUBANN_NOALLOC_SCRATCH_BYTES) {
      snprintf(msg, sizeof(msg),
               "no-alloc query: %u queries x %u table chunks x k=%u need %llu bytes of scratch, the fixed array holds %llu; "
               "send fewer queries per call or a smaller k",
               nq, cfg.nchunks, keff, (unsigned long long)need, (unsigned long long)JUBANN_NOALLOC_SCRATCH_BYTES);
      return set_error(JUBANN_E_INVALID, msg);
    }
    if ((nq + cfg.qp - 1) / cfg.qp > 65535u) return set_error(JUBANN_E_INVALID, "no-alloc query: too many query groups for one launch");
  }
  void* scratch = nullptr;
  rc = scratch_address(hi.device, &scratch);
  if (rc) return rc;

  cudaError_t e = cudaMemsetAsync(d_out_count, 0, (size_t)nq * 4, s);
  if (e != cudaSuccess) return cuda_error("cudaMemsetAsync(out_count)", e);
  if (keff == 0) return 0;
  uint64_t* tau_scratch = static_cast<uint64_t*>(scratch) + (size_t)nq * cfg.nchunks * keff;
  // euclid_lsh's two-level score histogram (257 KiB per query) does not fit this fixed array: that
  // engine runs here with each CTA's own bound only
  uint32_t* hist_scratch = hi.euclid ? nullptr : reinterpret_cast<uint32_t*>(tau_scratch + nq);
  int scan_kernels = 0;   // the scan, plus the sample pre-pass when it runs
  e = launch_scan(hi.table, hi.euclid, cfg, d_qsigs, d_qnorms, hi.coslut, nq, keff, nullptr, tau_scratch, hist_scratch,
                  static_cast<uint64_t*>(scratch), s, &scan_kernels);
  if (e != cudaSuccess) return cuda_error("scan launch (no-alloc)", e);
  MergeOut o{};
  o.kout = k; o.col0 = 0;
  o.shard_rank = 0; o.shard_count = 1;
  o.hash_num = hi.table.hash_num; o.euclid = hi.euclid ? 1 : 0; o.flags = flags;
  o.qnorms = d_qnorms;
  o.out_index = d_out_index; o.out_score = d_out_score; o.out_count = d_out_count;
  e = launch_merge_local(static_cast<const uint64_t*>(scratch), nq, cfg.nchunks, keff, o, s);
  if (e != cudaSuccess) return cuda_error("merge launch (no-alloc)", e);
  handle_add_launches(h, (uint64_t)scan_kernels + 1, 1, (nq + cfg.qp - 1) / cfg.qp);   // + the list reduction
  return 0;
}
