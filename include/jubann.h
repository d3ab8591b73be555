/*
 * jubann.h -- C-ABI of the B200-native nearest_neighbor backend (libjubann.so).
 *
 * The reference (Jubatus 0.9.2) has no C ABI and no plugin slot for NN engines: the
 * engine is a C++ object obtained from jubatus_core's factory at
 *   jubatus/server/server/nearest_neighbor_serv.cpp:106-108  (create_nearest_neighbor)
 * and the nested factories at recommender_serv.cpp:111-116 and anomaly_serv.cpp:134-138.
 * This library sits behind an adapter subclass of core::nearest_neighbor::
 * nearest_neighbor_base (jubatus_b200/host/) and is loaded the way the
 * reference loads its fv_converter plugins: dlopen(RTLD_LAZY), $JUBATUS_PLUGIN_PATH,
 * optional version() symbol (jubatus/server/fv_converter/dynamic_loader.cpp:44-94,
 * dynamic_loader.hpp:40-50).
 *
 * Conventions
 *  - Strings (row ids, feature keys) never cross this boundary.  The host owns
 *    key <-> global row index (column_table::index_/keys_) and the (owner, clock)
 *    versions; the library sees integers only.  A feature arrives as the 64-bit FNV-1
 *    hash of its key (common::hash_util::calc_string_hash) and/or its low 32 bits, which
 *    is what the engines seed their random vectors with.
 *  - Every function returns 0 on success or a negative JUBANN_E_* code; the message is
 *    in jubann_last_error() (thread-local).  Nothing aborts or exits.  CUDA errors are
 *    sticky on the handle (JUBANN_E_CUDA thereafter), like a jubatus_exception that the
 *    RPC dispatcher turns into an error string (common/mprpc/rpc_server.cpp:41-53).
 *  - Unless a name ends in _dev, all pointers are caller-owned HOST memory valid for the
 *    duration of the call; the call returns after the device work has completed.
 *    _dev entry points take DEVICE pointers plus the CUDA stream to enqueue on
 *    (cudaStream_t passed as void*, NULL = legacy default stream) and return without
 *    synchronising.
 *  - Thread safety: any number of threads may call concurrently on one handle (queries
 *    are NOLOCK_ in the reference, nearest_neighbor_impl.cpp:75-102); calls are
 *    serialised inside the library.
 *  - Row order: a row's *global row index* is its position in the reference's
 *    column_table (insertion order; delete_row moves the last row into the hole).  Ties
 *    in a ranking are broken by smaller global row index, exactly like the
 *    std::pair<dist, index> heap of ranking_hamming_bit_vectors.
 *  - Sharding: with shard_count G > 1 a handle stores only rows whose global index g
 *    satisfies g % G == shard_rank (local slot g / G); it may be handed every row (it keeps
 *    its own) or only its own.  Once the shards share a communicator (jubann_comm_init, or
 *    jubann_create_multi) the query entry points answer for the whole table; without one,
 *    jubann_scan_local_dev gives the shard-local top-k and jubann_merge_dev combines G such lists.
 *  - Ordering: host-pointer entry points run on a stream of the handle and return when the
 *    work is done.  _dev entry points enqueue on the caller's stream; work the handle does
 *    afterwards on its own stream (any host-pointer call, growing the table) is ordered after
 *    _dev writes by the library.  Two different caller streams are the caller's to order.
 *  - A batch that names one destination row more than once stores the last entry (what
 *    column_table::add does when applied in order).  Signature words are taken modulo
 *    hash_num: bits at or above it are ignored (bit_vector keeps them zero).
 */
#ifndef JUBANN_H_
#define JUBANN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct jubann jubann_t;

#define JUBANN_ABI_VERSION 2

#define JUBANN_OK 0
#define JUBANN_E_INVALID (-1)      /* invalid_parameter / unsupported_method */
#define JUBANN_E_CUDA (-2)         /* CUDA runtime error (sticky) */
#define JUBANN_E_NOMEM (-3)        /* device or host allocation failed */
#define JUBANN_E_CACHE_FULL (-4)   /* (ABI 1; no longer returned: features beyond the cache are drawn on the spot) */
#define JUBANN_E_RANGE (-5)        /* row index out of range */

/* query flags */
#define JUBANN_SIMILAR 1u  /* apply calc_similarity: 1-d (lsh, minhash), -d (euclid_lsh) */

/* Plugin-convention version symbol (dynamic_loader.hpp:40-50 looks up "version"). */
const char* version(void);
int jubann_abi_version(void);
const char* jubann_last_error(void);

/* Replaces core::nearest_neighbor::create_nearest_neighbor(method, param, table, my_id)
 * (call site nearest_neighbor_serv.cpp:106-108) together with `new column_table`
 * (:99-100): method is "lsh" | "euclid_lsh" | "minhash" (else JUBANN_E_INVALID, the
 * factory's unsupported_method); hash_num is parameter.hash_num (config/nearest_neighbor/
 * lsh.json:17-19; < 1 is invalid_parameter; this build supports 1..1024).
 * device = CUDA ordinal.  shard_rank/shard_count: see "Sharding" above (1 GPU: 0, 1). */
int jubann_create(const char* method, uint32_t hash_num, int device, uint32_t shard_rank,
                  uint32_t shard_count, jubann_t** out);
int jubann_destroy(jubann_t* h);

/* Tunables, to be set before the first row or query.  Names: "cache_slots" (distinct
 * feature seeds whose random vectors the projection cache keeps, default 2^20; features met
 * after it is full are never refused: their vectors are drawn on the spot, as
 * lsh_function::random_projection does for every feature), "queries_per_pass" (how many
 * queries share one streaming pass over the table; 0 = automatic), "scan_variant"
 * (0 = automatic, 1 = LDG streaming, 2 = TMA bulk-copy pipeline), "popc_variant" (0 = one
 * POPC per word, 1..4 = carry-save adder trees of growing depth, default 3), "scan_order" (1 = CTAs take turns over
 * consecutive tile groups so the whole GPU sweeps one front, 0 = one contiguous range per CTA),
 * "reserve_rows" (allocate the slab for this
 * many global rows up front; may be set at any time).  Two switches select alternates of choices
 * that jubatus_core's source would settle (0 = the frozen choice the parity tests
 * run under): "gaussian_form" (1: the Box-Muller variate is taken with sin instead of cos; before
 * the first row or query) and "euclid_score" (1: cosf and the whole score expression in float
 * instead of double). */
int jubann_set_option(jubann_t* h, const char* name, int64_t value);

/* nearest_neighbor_base::clear() -> column_table::clear()  (nearest_neighbor_serv.cpp:125) */
int jubann_clear(jubann_t* h);
/* column_table::size(): number of GLOBAL rows this handle has been told about
 * (max global index + 1 after truncation), and rows stored locally. */
int jubann_size(jubann_t* h, uint64_t* global_rows, uint64_t* local_rows);

/* Signature generation only (lsh_function::random_projection + binarize,
 * euclid_lsh::calc_norm, minhash::hash) for n sparse vectors in CSR form:
 * indptr[n+1] offsets into seed32/keyhash64/vals.  seed32[i] = (uint32_t)FNV1_64(key_i)
 * is required for lsh/euclid_lsh, keyhash64[i] = FNV1_64(key_i) for minhash (the other may
 * be NULL).  out_sigs: n * ceil(hash_num/64) uint64 words, bit j at word j/64 bit j%64
 * (storage::bit_vector layout); out_norms: n floats (0 unless euclid_lsh), may be NULL. */
int jubann_hash_csr(jubann_t* h, uint64_t n, const uint64_t* indptr, const uint32_t* seed32,
                    const uint64_t* keyhash64, const float* vals, uint64_t* out_sigs,
                    float* out_norms);

/* {lsh,minhash,euclid_lsh}::set_row -> column_table::add  (reached from
 * nearest_neighbor_serv.cpp:134): hash each sparse vector and store it at global row index
 * row_index[r] (append when row_index[r] == current global size, overwrite when smaller;
 * indices must not skip).  Rows not owned by this shard are ignored. */
int jubann_set_rows_csr(jubann_t* h, uint64_t n, const uint64_t* row_index, const uint64_t* indptr,
                        const uint32_t* seed32, const uint64_t* keyhash64, const float* vals);

/* column_table::set_row(packed) as used by mixable_versioned_table::push
 * (framework/mixer/push_mixer.cpp:449-472) and driver unpack (framework/save_load.cpp:280):
 * store given signatures (+ norms for euclid_lsh, else NULL). */
int jubann_set_rows_sig(jubann_t* h, uint64_t n, const uint64_t* row_index, const uint64_t* sigs,
                        const float* norms);
/* column_table::get_row as used by mixable_versioned_table::pull (push_mixer.cpp:410-431)
 * and driver pack (save_load.cpp:129).  Rows not owned by this shard are zero-filled. */
int jubann_get_rows_sig(jubann_t* h, uint64_t n, const uint64_t* row_index, uint64_t* out_sigs,
                        float* out_norms);

/* column_table::delete_row(key) = move the last row into the hole, shrink by one.  The
 * host calls move_row(last, hole) (when they differ) then truncate(size-1).  In sharded
 * mode every shard receives both calls; sigs/norms of `src` are fetched from whichever
 * shard owns it by the host (get_rows_sig) and stored with set_rows_sig, so move_row is
 * the single-shard fast path (both rows local). */
int jubann_move_row(jubann_t* h, uint64_t src_index, uint64_t dst_index);
int jubann_truncate(jubann_t* h, uint64_t new_global_rows);

/* bit_vector_nearest_neighbor_base::neighbor_row_from_hash -> ranking_hamming_bit_vectors
 * (lsh, minhash) / euclid_lsh::neighbor_row_from_hash; with JUBANN_SIMILAR also
 * nearest_neighbor_base::similar_row.  nq query signatures (same layout as out_sigs above;
 * qnorms only for euclid_lsh).  Outputs are [nq][k]: out_index = global row index,
 * out_score = dist/hash_num | sqrt(nq^2+score) | similarity; out_count[q] = min(k, rows). */
int jubann_query_sig(jubann_t* h, uint64_t nq, const uint64_t* qsigs, const float* qnorms,
                     uint32_t k, uint32_t flags, uint64_t* out_index, float* out_score,
                     uint32_t* out_count);
/* neighbor_row(sfv) / similar_row(sfv): hash the query datums, then as above
 * (neighbor_row_from_datum / similar_row_from_datum, nearest_neighbor_serv.cpp:147-172). */
int jubann_query_csr(jubann_t* h, uint64_t nq, const uint64_t* indptr, const uint32_t* seed32,
                     const uint64_t* keyhash64, const float* vals, uint32_t k, uint32_t flags,
                     uint64_t* out_index, float* out_score, uint32_t* out_count);
/* neighbor_row(id) / similar_row(id): the query is the STORED signature of a row
 * (no re-hash; nearest_neighbor_serv.cpp:138-145,156-163).  Single-shard handles only;
 * sharded callers fetch the signature with get_rows_sig and use query_sig. */
int jubann_query_rows(jubann_t* h, uint64_t nq, const uint64_t* row_index, uint32_t k,
                      uint32_t flags, uint64_t* out_index, float* out_score, uint32_t* out_count);

/* set_row immediately followed by neighbor_row(id) / similar_row(id) of the row just stored: what
 * anomaly add does (anomaly_serv.cpp:157-176 -> [core] light_lof::set_row, which stores the row and at
 * once collects its neighbours) and what set_row + neighbor_row_from_id give when issued back to back
 * (nearest_neighbor_serv.cpp:134,144).  The sparse vector is hashed once, stored at row_index (append or
 * overwrite, as jubann_set_rows_csr) and ranked against the table that now contains it: one host round
 * trip instead of two.  Single-shard handles; rows whose CSR does not fit the 1 MiB staging buffer take
 * the two separate calls. */
int jubann_set_row_query_csr(jubann_t* h, uint64_t row_index, uint64_t nnz, const uint32_t* seed32, const uint64_t* keyhash64,
                             const float* vals, uint32_t k, uint32_t flags, uint64_t* out_index, float* out_score,
                             uint32_t* out_count);

/* ---- device-pointer entry points (bench, torch.distributed / NCCL plumbing) ---- */

/* jubann_query_sig with device-resident queries and outputs; the library owns and grows the
 * scratch it needs.  Collective on a sharded handle with a communicator (see below). */
int jubann_query_sig_dev(jubann_t* h, uint64_t nq, const uint64_t* d_qsigs, const float* d_qnorms,
                         uint32_t k, uint32_t flags, uint64_t* d_out_index, float* d_out_score,
                         uint32_t* d_out_count, void* stream);
int jubann_set_rows_sig_dev(jubann_t* h, uint64_t n, uint64_t first_row_index,
                            const uint64_t* d_sigs, const float* d_norms, void* stream);

/* Device-resident forms of jubann_hash_csr / jubann_set_rows_csr (same reference functions:
 * lsh_function::random_projection + binarize, euclid_lsh::calc_norm, minhash::hash, then
 * column_table::add): the CSR batch is already in device memory (indptr[n+1] starting at 0,
 * nnz = indptr[n]); rows are stored at first_row_index .. first_row_index + n - 1.  For
 * lsh / euclid_lsh the stream is synchronised once per call (features never seen before get
 * their projection-cache rows first). */
int jubann_hash_csr_dev(jubann_t* h, uint64_t n, uint64_t nnz, const uint64_t* d_indptr, const uint32_t* d_seed32,
                        const uint64_t* d_keyhash64, const float* d_vals, uint64_t* d_out_sigs, float* d_out_norms,
                        void* stream);
int jubann_set_rows_csr_dev(jubann_t* h, uint64_t n, uint64_t first_row_index, uint64_t nnz, const uint64_t* d_indptr,
                            const uint32_t* d_seed32, const uint64_t* d_keyhash64, const float* d_vals, void* stream);

/* jubann_set_rows_csr_dev for rows first_row_index + i * row_stride: a shard of a table spread over
 * G ranks is bulk-loaded with its own rows only (row_stride = shard_count, first_row_index a row of
 * this shard), so that no rank hashes a row it does not keep.  Same reference functions
 * ({lsh,minhash,euclid_lsh}::set_row -> column_table::add, nearest_neighbor_serv.cpp:134) behind the
 * reference's placement of a row on its owner (#@cht(1), nearest_neighbor.idl:10). */
int jubann_set_rows_csr_strided_dev(jubann_t* h, uint64_t n, uint64_t first_row_index, uint64_t row_stride, uint64_t nnz,
                                    const uint64_t* d_indptr, const uint32_t* d_seed32, const uint64_t* d_keyhash64,
                                    const float* d_vals, void* stream);

/* No counterpart in the reference.  Same inputs, outputs and results as
 * jubann_query_sig_dev, for callers that may not allocate device memory once they are
 * running: a server whose GPU memory is parcelled out at start-up by its own pool, RPC
 * threads with a latency budget (cudaMalloc/cudaFree stall the whole device), a process in
 * which another thread may be recording work on a stream.  The kernels' scratch (per-CTA
 * candidate lists) is a fixed array of JUBANN_NOALLOC_SCRATCH_BYTES inside the library image
 * instead of a buffer grown with cudaMalloc; every GPU that loads the library holds it until
 * exit.  This call allocates and frees nothing.  It needs nq * (chunks * min(k, rows) + 1) * 8
 * + nq * (hash_num + 33) * 4 bytes (that term for lsh / minhash only; chunks <= 2 per SM); a larger request, or k above 2048,
 * returns JUBANN_E_INVALID
 * with a message.  The first call on a device looks up the array's address
 * (cudaGetSymbolAddress): make it at start-up, before any stream capture.  Calls share the
 * array and it keeps the previous call's contents (the part that must start at zero, a small
 * histogram, is cleared on the stream by each call); calls whose GPU work overlaps race on it --
 * keep them on one stream.  Single-shard
 * handles only. */
#define JUBANN_NOALLOC_SCRATCH_BYTES (32u << 20)
int jubann_query_sig_noalloc_dev(jubann_t* h, uint64_t nq, const uint64_t* d_qsigs, const float* d_qnorms,
                                 uint32_t k, uint32_t flags, uint64_t* d_out_index, float* d_out_score,
                                 uint32_t* d_out_count, void* stream);

/* Sharded query, stage 1: shard-local top-k as raw candidates, [nq][k] each:
 * d_cand_key = ordering key (Hamming distance, or the order-preserving image of the
 * pre-sqrt euclid score), d_cand_index = GLOBAL row index; unused slots hold
 * (0xFFFFFFFF, UINT64_MAX).  These are what the ranks all-gather (NCCL over NVLink). */
int jubann_scan_local_dev(jubann_t* h, uint64_t nq, const uint64_t* d_qsigs, const float* d_qnorms,
                          uint32_t k, uint32_t* d_cand_key, uint64_t* d_cand_index, void* stream);
/* Sharded query, stage 2: merge `parts` gathered candidate lists laid out [parts][nq][k]
 * by (key, global index) ascending -- the same total order as the single-table heap --
 * and produce the final [nq][k] outputs (scores as in jubann_query_sig). */
int jubann_merge_dev(jubann_t* h, uint64_t nq, uint32_t k, uint32_t parts, const uint32_t* d_cand_key,
                     const uint64_t* d_cand_index, const float* d_qnorms, uint32_t flags,
                     uint64_t* d_out_index, float* d_out_score, uint32_t* d_out_count, void* stream);

/* MIX (mixable_versioned_table get_diff/put_diff as an all-gather of row diffs,
 * linear_mixer.cpp:547-561,613-628): pack the listed rows into fixed-size records
 * [row_index u64 | sig words | norm f32 padded to 8 B] in device memory, and apply such
 * records.  The version arbitration (newer clock wins) is the host's; see INTEGRATION.md. */
int jubann_diff_record_bytes(jubann_t* h, uint32_t* bytes);
int jubann_diff_pack_dev(jubann_t* h, uint64_t n, const uint64_t* d_row_index, void* d_records,
                         void* stream);
int jubann_diff_apply_dev(jubann_t* h, uint64_t n, const void* d_records, void* stream);

/* ---- more than one GPU (NCCL over NVLink under this library) -------------------------------
 *
 * The reference spreads one model over servers with its proxy (framework/proxy.hpp:249-266
 * broadcasts a query, aggregators.hpp:27-63 fold the answers) and keeps replicas in step with
 * its mixers (framework/mixer/push_mixer.cpp:410-472, linear_mixer.cpp:547-561,613-628).  Here
 * the GPUs of one box play those roles and the exchanges are collectives inside the library, so
 * that the C++ host needs nothing but these calls.
 *
 * (1) One process, several GPUs -- SURVEY 8(b)'s jubann_create(..., devices[], ndev, ...):
 * replaces create_nearest_neighbor + new column_table (nearest_neighbor_serv.cpp:98-109) by a
 * table whose rows are sharded round-robin over devices[0..ndev).  The handle answers every
 * host-pointer entry point above (set_option, clear, size, hash_csr, set_rows_*, get_rows_sig,
 * move_row, truncate, query_*, stats); a query is: signatures to every GPU, shard-local scan,
 * one all-gather of packed (key, global index) candidates, merge -- bit-exactly the single-table
 * answer.  _dev entry points refuse such a handle. */
int jubann_create_multi(const char* method, uint32_t hash_num, const int* devices, int ndev, jubann_t** out);

/* (2) One process per GPU.  Rank 0 obtains an id (JUBANN_COMM_ID_BYTES bytes) and hands it to the
 * other ranks by whatever transport the host has (the Jubatus servers: ZooKeeper, where they
 * already register themselves -- common/membership.cpp; torchrun: the store); then every rank
 * joins with its handle.  Collective: returns when all `world` ranks have called.  For a handle
 * created with shard_rank/shard_count the communicator must have the same rank/world; from then
 * on jubann_query_sig, jubann_query_csr and jubann_query_sig_dev are collective on such a handle:
 * every rank makes the same call with the same queries and receives the same, global answer.
 * Handles of replicas (shard_count 1) use the communicator for MIX. */
#define JUBANN_COMM_ID_BYTES 128
int jubann_comm_unique_id(void* id);
int jubann_comm_init(jubann_t* h, const void* id, uint32_t rank, uint32_t world);
int jubann_comm_info(jubann_t* h, uint32_t* rank, uint32_t* world);

/* MIX between replicas as collectives (mixable_versioned_table get_diff / put_diff,
 * linear_mixer.cpp:547-561,613-628; pull / push, push_mixer.cpp:410-472).  Every replica calls
 * jubann_mix_allgather with the rows it has changed since the last round (row_index: n local
 * rows) and an opaque blob the host has serialised for them (keys, owners, clocks: strings stay
 * the host's business).  One all-gather of the counts and one of the payload -- the rows'
 * signature records packed on the device, followed by the blob -- leave on every replica:
 * out_rows[r] / out_blob_bytes[r] for each rank r, out_blobs[r] = rank r's blob (pinned memory
 * owned by the handle, valid until the next MIX call), and the gathered records on the device.
 * The host then decides, per gathered record, whether it is newer than what it holds (newer
 * clock wins; unknown keys append) and calls jubann_mix_apply with the m records to store:
 * record src_pos[e] of rank src_rank[e] goes to local row dst_row[e] (append when equal to the
 * current size; if a row is named twice the later entry wins). */
int jubann_mix_allgather(jubann_t* h, uint64_t n, const uint64_t* row_index, const void* blob, uint64_t blob_bytes,
                         uint64_t* out_rows, uint64_t* out_blob_bytes, const void** out_blobs);
int jubann_mix_apply(jubann_t* h, uint64_t m, const uint32_t* src_rank, const uint32_t* src_pos,
                     const uint64_t* dst_row);

/* Introspection for bench.py / tests: kernels launched by this handle since the last reset;
 * scan-kernel launches; streaming passes over the table (a launch covers several passes
 * when the batch holds more queries than share one pass); and, when switched on with
 * jubann_stats_reset(h, 1), the summed duration of the scan kernels in ms (CUDA events
 * on the launching stream). */
int jubann_stats(jubann_t* h, uint64_t* kernel_launches, uint64_t* scan_launches, uint64_t* scan_passes,
                 double* scan_ms);
int jubann_stats_reset(jubann_t* h, int time_scan_kernels);

#ifdef __cplusplus
}
#endif
#endif /* JUBANN_H_ */
