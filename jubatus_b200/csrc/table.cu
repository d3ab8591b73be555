// Device-resident column_table: signature slab (+ norm column) mutation kernels (sm_100a).
//
// Replaces the storage side of [core] storage/column_table.hpp (add / update / set_row /
// get_row / delete_row's move-last-into-hole) and storage/bit_vector.hpp for the columns the
// nearest_neighbor engines use; constructed by the reference at
// jubatus/server/server/nearest_neighbor_serv.cpp:99-100.  Keys, the key->index map and the
// (owner, clock) versions stay on the host (jubatus_b200/host/).
#include "internal.h"

namespace jbn {

// one thread per (row, 32-bit word): row-major ABI words -> tile-transposed slab
__global__ void scatter_rows_kernel(TableView t, uint64_t n, const uint64_t* __restrict__ row_index,
                                    uint64_t first_index, uint64_t stride, const uint64_t* __restrict__ sigs,
                                    const float* __restrict__ norms, uint32_t rank, uint32_t G) {
  const uint64_t total = n * t.w32;
  unsigned char* slab = static_cast<unsigned char*>(t.slab);
  for (uint64_t e = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; e < total;
       e += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t r = e / t.w32;
    const uint32_t w = (uint32_t)(e % t.w32);
    const uint64_t g = row_index ? row_index[r] : first_index + r * stride;
    // IDX_SENTINEL: an entry the host struck out (a later entry of the batch writes the same row)
    if (g == IDX_SENTINEL || g % G != rank) continue;
    const uint64_t i = g / G;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(sigs + r * t.w64);
    const uint32_t v = (w < 2 * t.w64) ? (src[w] & sig_word_mask(t.hash_num, w)) : 0u;
    *reinterpret_cast<uint32_t*>(slab + word_offset(t.w32, i, w)) = v;
    if (w == 0 && t.norms) t.norms[i] = norms ? norms[r] : 0.0f;
  }
}

__global__ void gather_rows_kernel(TableView t, uint64_t n, const uint64_t* __restrict__ row_index,
                                   uint64_t* __restrict__ sigs, float* __restrict__ norms, uint32_t rank,
                                   uint32_t G) {
  const uint32_t aw = 2 * t.w64;
  const uint64_t total = n * aw;
  const unsigned char* slab = static_cast<const unsigned char*>(t.slab);
  for (uint64_t e = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; e < total;
       e += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t r = e / aw;
    const uint32_t w = (uint32_t)(e % aw);
    const uint64_t g = row_index[r];
    const bool mine = (g % G == rank) && (g / G < t.local_rows);
    const uint64_t i = g / G;
    uint32_t* dst = reinterpret_cast<uint32_t*>(sigs + r * t.w64);
    dst[w] = mine ? *reinterpret_cast<const uint32_t*>(slab + word_offset(t.w32, i, w)) : 0u;
    if (w == 0 && norms) norms[r] = (mine && t.norms) ? t.norms[i] : 0.0f;
  }
}

// record = [row_index u64][w64 sig words u64][norm f32][pad u32]
__global__ void diff_pack_kernel(TableView t, uint64_t n, const uint64_t* __restrict__ row_index,
                                 unsigned char* __restrict__ rec, uint32_t rec_bytes, uint32_t rank, uint32_t G) {
  const uint32_t aw = 2 * t.w64;
  const uint64_t total = n * (aw + 1);
  const unsigned char* slab = static_cast<const unsigned char*>(t.slab);
  for (uint64_t e = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; e < total;
       e += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t r = e / (aw + 1);
    const uint32_t w = (uint32_t)(e % (aw + 1));
    const uint64_t g = row_index[r];
    const bool mine = (g % G == rank) && (g / G < t.local_rows);
    const uint64_t i = g / G;
    unsigned char* out = rec + r * rec_bytes;
    if (w == aw) {
      *reinterpret_cast<uint64_t*>(out) = mine ? g : IDX_SENTINEL;
      *reinterpret_cast<float*>(out + 8 + 8ull * t.w64) = (mine && t.norms) ? t.norms[i] : 0.0f;
      *reinterpret_cast<uint32_t*>(out + 12 + 8ull * t.w64) = 0u;
    } else {
      reinterpret_cast<uint32_t*>(out + 8)[w] =
          mine ? *reinterpret_cast<const uint32_t*>(slab + word_offset(t.w32, i, w)) : 0u;
    }
  }
}

__global__ void diff_apply_kernel(TableView t, uint64_t n, const unsigned char* __restrict__ rec,
                                  uint32_t rec_bytes, uint32_t rank, uint32_t G) {
  const uint64_t total = n * t.w32;
  unsigned char* slab = static_cast<unsigned char*>(t.slab);
  for (uint64_t e = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; e < total;
       e += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t r = e / t.w32;
    const uint32_t w = (uint32_t)(e % t.w32);
    const unsigned char* in = rec + r * rec_bytes;
    const uint64_t g = *reinterpret_cast<const uint64_t*>(in);
    if (g == IDX_SENTINEL || g % G != rank) continue;
    const uint64_t i = g / G;
    if (i >= t.local_rows) continue;
    const uint32_t v = (w < 2 * t.w64) ? (reinterpret_cast<const uint32_t*>(in + 8)[w] & sig_word_mask(t.hash_num, w)) : 0u;
    *reinterpret_cast<uint32_t*>(slab + word_offset(t.w32, i, w)) = v;
    if (w == 0 && t.norms) t.norms[i] = *reinterpret_cast<const float*>(in + 8 + 8ull * t.w64);
  }
}

// MIX: entry e copies record src_pos[e] of rank src_rank[e]'s slot of the all-gathered payload
// into local row dst_row[e] (IDX_SENTINEL = struck out by the host: an older version, or a later
// entry of this round writes the same row)
__global__ void mix_apply_kernel(TableView t, uint64_t m, const unsigned char* __restrict__ gathered, uint64_t slot_bytes,
                                 uint32_t rec_bytes, const uint32_t* __restrict__ src_rank,
                                 const uint32_t* __restrict__ src_pos, const uint64_t* __restrict__ dst_row) {
  const uint64_t total = m * t.w32;
  unsigned char* slab = static_cast<unsigned char*>(t.slab);
  for (uint64_t e = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; e < total;
       e += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t r = e / t.w32;
    const uint32_t w = (uint32_t)(e % t.w32);
    const uint64_t i = dst_row[r];
    if (i == IDX_SENTINEL || i >= t.local_rows) continue;
    const unsigned char* in = gathered + (uint64_t)src_rank[r] * slot_bytes + (uint64_t)src_pos[r] * rec_bytes;
    const uint32_t v = (w < 2 * t.w64) ? (reinterpret_cast<const uint32_t*>(in + 8)[w] & sig_word_mask(t.hash_num, w)) : 0u;
    *reinterpret_cast<uint32_t*>(slab + word_offset(t.w32, i, w)) = v;
    if (w == 0 && t.norms) t.norms[i] = *reinterpret_cast<const float*>(in + 8 + 8ull * t.w64);
  }
}

static unsigned grid_for(uint64_t total, int num_sms) {
  uint64_t b = (total + 255) / 256;
  if (b > (uint64_t)num_sms * 32) b = (uint64_t)num_sms * 32;
  if (b < 1) b = 1;
  return (unsigned)b;
}

cudaError_t launch_scatter_rows(const TableView& t, uint64_t n, const uint64_t* d_row_index,
                                uint64_t first_index, const uint64_t* d_sigs, const float* d_norms,
                                uint32_t shard_rank, uint32_t shard_count, int num_sms, cudaStream_t s,
                                uint64_t index_stride) {
  if (n == 0) return cudaSuccess;
  scatter_rows_kernel<<<grid_for(n * t.w32, num_sms), 256, 0, s>>>(t, n, d_row_index, first_index, index_stride, d_sigs,
                                                                   d_norms, shard_rank, shard_count);
  return cudaGetLastError();
}
cudaError_t launch_gather_rows(const TableView& t, uint64_t n, const uint64_t* d_row_index,
                               uint64_t* d_sigs, float* d_norms, uint32_t shard_rank,
                               uint32_t shard_count, int num_sms, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  gather_rows_kernel<<<grid_for(n * 2 * t.w64, num_sms), 256, 0, s>>>(t, n, d_row_index, d_sigs, d_norms, shard_rank,
                                                             shard_count);
  return cudaGetLastError();
}
cudaError_t launch_diff_pack(const TableView& t, uint64_t n, const uint64_t* d_row_index, void* d_records,
                             uint32_t record_bytes, uint32_t shard_rank, uint32_t shard_count, int num_sms,
                             cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  diff_pack_kernel<<<grid_for(n * (2 * t.w64 + 1), num_sms), 256, 0, s>>>(
      t, n, d_row_index, static_cast<unsigned char*>(d_records), record_bytes, shard_rank, shard_count);
  return cudaGetLastError();
}
cudaError_t launch_diff_apply(const TableView& t, uint64_t n, const void* d_records, uint32_t record_bytes,
                              uint32_t shard_rank, uint32_t shard_count, int num_sms, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  diff_apply_kernel<<<grid_for(n * t.w32, num_sms), 256, 0, s>>>(t, n, static_cast<const unsigned char*>(d_records),
                                                        record_bytes, shard_rank, shard_count);
  return cudaGetLastError();
}

cudaError_t launch_mix_apply(const TableView& t, uint64_t m, const void* d_gathered, uint64_t slot_bytes,
                             uint32_t record_bytes, const uint32_t* d_src_rank, const uint32_t* d_src_pos,
                             const uint64_t* d_dst_row, int num_sms, cudaStream_t s) {
  if (m == 0) return cudaSuccess;
  mix_apply_kernel<<<grid_for(m * t.w32, num_sms), 256, 0, s>>>(t, m, static_cast<const unsigned char*>(d_gathered), slot_bytes,
                                                               record_bytes, d_src_rank, d_src_pos, d_dst_row);
  return cudaGetLastError();
}

}  // namespace jbn
