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
                                    uint64_t first_index, const uint64_t* __restrict__ sigs,
                                    const float* __restrict__ norms, uint32_t rank, uint32_t G,
                                    const uint32_t* skip_if_nonzero) {
  // latency path: the projection cache overflowed in this very call, the signatures are not valid
  if (skip_if_nonzero && *reinterpret_cast<const volatile uint32_t*>(skip_if_nonzero)) return;
  const uint64_t total = n * t.w32;
  unsigned char* slab = static_cast<unsigned char*>(t.slab);
  for (uint64_t e = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; e < total;
       e += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t r = e / t.w32;
    const uint32_t w = (uint32_t)(e % t.w32);
    const uint64_t g = row_index ? row_index[r] : first_index + r;
    if (g % G != rank) continue;
    const uint64_t i = g / G;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(sigs + r * t.w64);
    const uint32_t v = (w < 2 * t.w64) ? src[w] : 0u;
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
    const uint32_t v = (w < 2 * t.w64) ? reinterpret_cast<const uint32_t*>(in + 8)[w] : 0u;
    *reinterpret_cast<uint32_t*>(slab + word_offset(t.w32, i, w)) = v;
    if (w == 0 && t.norms) t.norms[i] = *reinterpret_cast<const float*>(in + 8 + 8ull * t.w64);
  }
}

static unsigned grid_for(uint64_t total) {
  uint64_t b = (total + 255) / 256;
  if (b > 148ull * 32) b = 148ull * 32;
  if (b < 1) b = 1;
  return (unsigned)b;
}

cudaError_t launch_scatter_rows(const TableView& t, uint64_t n, const uint64_t* d_row_index,
                                uint64_t first_index, const uint64_t* d_sigs, const float* d_norms,
                                uint32_t shard_rank, uint32_t shard_count, cudaStream_t s,
                                const uint32_t* d_skip_if_nonzero) {
  if (n == 0) return cudaSuccess;
  scatter_rows_kernel<<<grid_for(n * t.w32), 256, 0, s>>>(t, n, d_row_index, first_index, d_sigs, d_norms,
                                                          shard_rank, shard_count, d_skip_if_nonzero);
  return cudaGetLastError();
}
cudaError_t launch_gather_rows(const TableView& t, uint64_t n, const uint64_t* d_row_index,
                               uint64_t* d_sigs, float* d_norms, uint32_t shard_rank,
                               uint32_t shard_count, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  gather_rows_kernel<<<grid_for(n * 2 * t.w64), 256, 0, s>>>(t, n, d_row_index, d_sigs, d_norms, shard_rank,
                                                             shard_count);
  return cudaGetLastError();
}
cudaError_t launch_diff_pack(const TableView& t, uint64_t n, const uint64_t* d_row_index, void* d_records,
                             uint32_t record_bytes, uint32_t shard_rank, uint32_t shard_count,
                             cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  diff_pack_kernel<<<grid_for(n * (2 * t.w64 + 1)), 256, 0, s>>>(
      t, n, d_row_index, static_cast<unsigned char*>(d_records), record_bytes, shard_rank, shard_count);
  return cudaGetLastError();
}
cudaError_t launch_diff_apply(const TableView& t, uint64_t n, const void* d_records, uint32_t record_bytes,
                              uint32_t shard_rank, uint32_t shard_count, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  diff_apply_kernel<<<grid_for(n * t.w32), 256, 0, s>>>(t, n, static_cast<const unsigned char*>(d_records),
                                                        record_bytes, shard_rank, shard_count);
  return cudaGetLastError();
}

}  // namespace jbn
