// Internal declarations shared by the translation units of libjubann.so.
// Device-memory layout and kernel contracts are described in DESIGN.md.
#ifndef JUBANN_INTERNAL_H_
#define JUBANN_INTERNAL_H_

#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <string>
#include <vector>

namespace jbn {

enum Method { LSH = 0, EUCLID_LSH = 1, MINHASH = 2 };

static const uint32_t KEY_SENTINEL_HI = 0xFFFFFFFFu;
static const uint64_t KEY_SENTINEL = 0xFFFFFFFFFFFFFFFFull;
static const uint64_t IDX_SENTINEL = 0xFFFFFFFFFFFFFFFFull;

// ---- signature slab layout -------------------------------------------------
// Rows are grouped in tiles of 32.  A row holds W32 32-bit words (hash_num padded up to
// 64, 128, 256, 512 or 1024 bits), cut into chunks of V bytes (V = 8 when W32 == 2, else
// 16).  Chunk c of the 32 rows of tile t is contiguous:
//     byte offset = ((t * C + c) * 32 + lane) * V,   C = W32 * 4 / V
// so one warp instruction (lane = row in tile) reads 32*V contiguous bytes.
__host__ __device__ inline uint32_t padded_w32(uint32_t hash_num) {
  uint32_t w = 2;
  while (w * 32 < hash_num) w <<= 1;
  return w;
}
__host__ __device__ inline uint32_t vec_bytes(uint32_t w32) { return w32 == 2 ? 8u : 16u; }
__host__ __device__ inline uint64_t slab_bytes(uint32_t w32, uint64_t rows) {
  return ((rows + 31) / 32) * 32ull * w32 * 4ull;
}
// byte offset of 32-bit word `w` of local row `i`
__host__ __device__ inline uint64_t word_offset(uint32_t w32, uint64_t i, uint32_t w) {
  const uint32_t V = vec_bytes(w32), C = w32 * 4 / V, wpc = V / 4;
  const uint64_t t = i >> 5;
  const uint32_t lane = (uint32_t)(i & 31), c = w / wpc, r = w % wpc;
  return ((t * C + c) * 32 + lane) * (uint64_t)V + r * 4ull;
}

// Bits of 32-bit word `w` of a signature that lie below hash_num.  Words handed in by callers
// (stored rows, MIX records, query signatures) are masked with it: storage::bit_vector keeps the
// unused high bits of its last word zero, and a stray bit there would push a Hamming distance
// past hash_num -- which indexes the cos table, the euclid bound table and the histogram bins.
__host__ __device__ inline uint32_t sig_word_mask(uint32_t hash_num, uint32_t w) {
  const uint32_t lo = w * 32u;
  if (lo >= hash_num) return 0u;
  const uint32_t left = hash_num - lo;
  return left >= 32u ? 0xFFFFFFFFu : ((1u << left) - 1u);
}

struct TableView {
  void* slab;          // tile-transposed signatures
  float* norms;        // [capacity] (euclid_lsh) or nullptr
  uint64_t local_rows;
  uint32_t w32;        // padded 32-bit words per row
  uint32_t w64;        // ABI words per row = ceil(hash_num / 64)
  uint32_t hash_num;
};

// ---- scan + top-k ----------------------------------------------------------
struct ScanConfig {
  int variant;        // 1 = LDG, 2 = TMA bulk pipeline
  int popc_variant;   // 0 plain, 1 one CSA level, 2 deep CSA tree
  uint32_t qp;        // queries per pass (CTA handles qp queries over its row chunk)
  uint32_t nchunks;   // grid.x
  uint32_t tiles_per_chunk;
  uint32_t cap;       // candidate list capacity per (CTA, query), power of two
  uint32_t stages, tiles_per_warp;  // TMA ring geometry (variant 2)
  uint32_t interleave;
  uint32_t score_mode;   // euclid_lsh: 0 = cos and score in double (frozen choice), 1 = cosf and score in float
  size_t smem_bytes;
};

// Chooses grid/smem for (table, nq, k).  Returns false if k does not fit one pass.
bool plan_scan(const TableView& t, bool euclid, uint32_t nq, uint32_t k, int variant, int popc_variant,
               uint32_t qp_opt, int interleave, int num_sms, ScanConfig* cfg);
uint32_t max_k_per_pass();
// d_hist_scratch (nullable, hamming engines): hist_scratch_bytes(nq, hash_num) of scratch for the
// distance histogram all CTAs share to derive a GPU-wide bound
uint32_t hist_stride_for(uint32_t hash_num, bool euclid);
size_t hist_scratch_bytes(uint32_t nq, uint32_t hash_num, bool euclid);
void set_tma_geometry_override(uint32_t tiles_per_warp, uint32_t stages);  // tuning experiments; 0 = automatic

// cand: [nq][nchunks][k] keys ((key_hi << 32) | local row) in no particular order, KEY_SENTINEL padded.
// floor_keys (nullable): per query exclusive lower bound (pagination when k > max_k_per_pass).
// d_tau_scratch (nullable): nq u64 of scratch for the sample pre-pass that seeds the thresholds.
cudaError_t launch_scan(const TableView& t, bool euclid, const ScanConfig& cfg, const uint64_t* d_qsigs,
                        const float* d_qnorms, const double* d_coslut, uint32_t nq, uint32_t k,
                        const uint64_t* d_floor_keys, uint64_t* d_tau_scratch, uint32_t* d_hist_scratch,
                        uint64_t* d_cand, cudaStream_t s, int* kernels_launched = nullptr);

// Per query: k smallest keys among nlists*k candidates.  If d_out_index != nullptr writes the
// final (index, score, count) outputs at column offset `col0` of rows of width `kout`;
// if d_cand_key != nullptr writes raw candidates (key_hi, global index).
struct MergeOut {
  uint64_t* out_index; float* out_score; uint32_t* out_count;   // final
  uint32_t* cand_key; uint64_t* cand_index;                      // raw
  uint64_t* last_key;                                            // [nq] k-th key (pagination), nullable
  uint32_t kout, col0;
  uint32_t shard_rank, shard_count;
  uint32_t hash_num; int euclid; uint32_t flags;
  const float* qnorms;
  int count_accumulate;  // add to out_count instead of overwriting (pagination)
};
// d_dense (nullable): nq words of scratch; with it the reduction first collects the few keys the scan
// left below its final thresholds and only falls back to selection for queries with thousands of them
cudaError_t launch_merge_local(const uint64_t* d_cand, uint32_t nq, uint32_t nlists, uint32_t k,
                               const MergeOut& out, cudaStream_t s, uint32_t* d_dense = nullptr,
                               int* kernels_launched = nullptr);
cudaError_t launch_merge_parts(const uint32_t* d_cand_key, const uint64_t* d_cand_index, uint32_t nq,
                               uint32_t k, uint32_t parts, const MergeOut& out, cudaStream_t s,
                               size_t key_stride = 0, size_t idx_stride = 0);   // strides in elements; 0 = nq * k

// ---- table mutation --------------------------------------------------------
// rows given row-major ([n][w64] u64 words) -> slab at local slot (g / G) when g % G == rank
cudaError_t launch_scatter_rows(const TableView& t, uint64_t n, const uint64_t* d_row_index,
                                uint64_t first_index, const uint64_t* d_sigs, const float* d_norms,
                                uint32_t shard_rank, uint32_t shard_count, int num_sms, cudaStream_t s,
                                uint64_t index_stride = 1);   // without d_row_index: row r goes to first_index + r * index_stride
cudaError_t launch_gather_rows(const TableView& t, uint64_t n, const uint64_t* d_row_index,
                               uint64_t* d_sigs, float* d_norms, uint32_t shard_rank,
                               uint32_t shard_count, int num_sms, cudaStream_t s);
cudaError_t launch_diff_pack(const TableView& t, uint64_t n, const uint64_t* d_row_index, void* d_records,
                             uint32_t record_bytes, uint32_t shard_rank, uint32_t shard_count, int num_sms,
                             cudaStream_t s);
cudaError_t launch_diff_apply(const TableView& t, uint64_t n, const void* d_records, uint32_t record_bytes,
                              uint32_t shard_rank, uint32_t shard_count, int num_sms, cudaStream_t s);

cudaError_t launch_mix_apply(const TableView& t, uint64_t m, const void* d_gathered, uint64_t slot_bytes,
                             uint32_t record_bytes, const uint32_t* d_src_rank, const uint32_t* d_src_pos,
                             const uint64_t* d_dst_row, int num_sms, cudaStream_t s);

// ---- signatures ------------------------------------------------------------
struct ProjCache {
  float* table;        // [slots][hash_num] gaussians narrowed to float
  uint64_t* keys;      // open-addressing keys (0 = empty, else seed | 1<<32)
  uint32_t* slot_of;   // slot per hash-table position
  uint32_t* counters;  // [0] slots handed out (may run past `slots`), [1] fill-list length, [2] feature occurrences met while full (statistic)
  uint2* fill_list;    // (slot, seed) pairs awaiting generation
  uint32_t slots;      // capacity
  uint32_t table_size; // hash table positions (power of two)
  uint32_t hash_num;
  uint32_t gauss_form; // 0: sqrt(-2 log(1-a)) cos(2 pi b) (frozen choice); 1: ... sin(2 pi b)
};
cudaError_t launch_cache_resolve(const ProjCache& c, const uint32_t* d_seed32, uint64_t nnz, int num_sms,
                                 cudaStream_t s, int* launches);
cudaError_t launch_cache_fill(const ProjCache& c, uint32_t nfill, int num_sms, cudaStream_t s, int* launches);
cudaError_t launch_cache_fill_queued(const ProjCache& c, uint64_t max_entries, int num_sms, cudaStream_t s, int* launches);
cudaError_t launch_signature_lsh(const ProjCache& c, bool euclid, uint64_t n, const uint64_t* d_indptr,
                                 const uint32_t* d_seed32, const float* d_vals, uint64_t* d_sigs,
                                 float* d_norms, uint32_t w64, int num_sms, bool may_miss, cudaStream_t s);
cudaError_t launch_signature_minhash(uint32_t hash_num, uint64_t n, const uint64_t* d_indptr,
                                     const uint64_t* d_keyhash64, const float* d_vals, uint64_t* d_sigs,
                                     float* d_norms, uint32_t w64, int num_sms, cudaStream_t s);


// ---- read-only view of a handle for translation units other than capi.cu ------------
struct HandleInfo {
  TableView table;
  int device;
  bool euclid;
  uint32_t shard_rank, shard_count;
  int variant, popc_variant, interleave;
  uint32_t qp;
  int num_sms;
  const double* coslut;
  int score_mode;
  std::mutex* mu;
};
}  // namespace jbn
struct jubann;
namespace jbn {
// fills `out`; returns 0, or a negative JUBANN_E_* code with the message set
int handle_info(jubann* h, HandleInfo* out);
void handle_add_launches(jubann* h, uint64_t kernels, uint64_t scans, uint64_t passes);
int set_error(int code, const char* message);
}  // namespace jbn
#endif
