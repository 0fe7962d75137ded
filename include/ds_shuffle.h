/*
 * ds_shuffle.h -- C ABI of the B200 (sm_100a) Datasets shuffle/sort hot path.
 *
 * Drop-in boundary for the per-block compute of the reference's Exoshuffle path
 * (Dataset.sort / random_shuffle / groupby().aggregate()).  In the reference this
 * arithmetic is done by pyarrow / numpy calls made from
 * python/ray/data/_internal/{arrow_block,sort,shuffle_and_partition}.py and
 * python/ray/data/_internal/arrow_ops/transform_pyarrow.py; each entry point below
 * cites the call site(s) it replaces (paths relative to python/ray/data/).
 *
 * Conventions
 *   - plain pointers + sizes; every pointer is a DEVICE pointer unless it is
 *     documented as host; no C++ or torch types cross this boundary.
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*),
 *     except where "host sync" is stated.
 *   - the library never allocates or frees device memory: temporaries come from
 *     a caller-provided workspace.  Two-phase protocol (like CUB): call with
 *     ws == NULL to get the required size in *ws_bytes, then call again.
 *   - the caller owns every buffer; nothing is retained after the call's work on
 *     `stream` completes.
 *   - return value: DS_OK (0) or a negative ds_status; ds_last_error() gives a
 *     thread-local message for the last failing call on this thread.
 *   - row counts are limited to DS_MAX_ROWS per call (32-bit row indices and
 *     30-bit look-back counters).
 */
#ifndef DS_SHUFFLE_H_
#define DS_SHUFFLE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int32_t ds_status;
#define DS_OK 0
#define DS_ERR_INVALID_ARG (-1)
#define DS_ERR_UNSUPPORTED (-2)
#define DS_ERR_WORKSPACE (-3)
#define DS_ERR_CUDA (-4)
#define DS_ERR_TOO_LARGE (-5)

#define DS_MAX_ROWS ((uint64_t)((1u << 30) - 1))
#define DS_MAX_KEY_BYTES 10 /* fixed_size_binary sort keys: 1..10 bytes */
#define DS_MAX_PARTITIONS 256

/* ds_sort_indices_* option bits */
#define DS_SORT_NO_PASS_SKIP 1 /* never read the digit histogram back: no host sync at all
                                  (implies DS_SORT_NO_PREFIX_REPAIR) */
#define DS_SORT_NO_PREFIX_REPAIR 2 /* always run every (non-constant) digit pass */
#define DS_SORT_NO_HYBRID 4 /* 10-byte keys: never finish the sort with one CTA per 16-bit bucket */

/* aggregation ops for ds_segmented_reduce_* */
#define DS_OP_SUM 0
#define DS_OP_MIN 1
#define DS_OP_MAX 2
#define DS_OP_ABSMAX 3 /* max of |x| (aggregate.py:289-305) */

/* Library version (major*10000 + minor*100 + patch).  No reference counterpart. */
int32_t ds_version(void);

/* Thread-local message of the last failing call.  Replaces the Python exceptions
 * the reference raises from these code paths (e.g. _internal/arrow_block.py:356-359). */
const char* ds_last_error(void);

/* ---------------------------------------------------------------- sort ---- */

/* Stable argsort of an int64 / uint64 / float64 key column.
 * Replaces pyarrow.compute.sort_indices(table, sort_keys=[(col, order)]) at
 * _internal/arrow_ops/transform_pyarrow.py:15 and :74 (called from
 * ArrowBlockAccessor.sort_and_partition, _internal/arrow_block.py:366-369, and
 * merge_sorted_blocks, _internal/arrow_block.py:473-486).
 *   col            n keys
 *   descending     0: ascending; 1: Arrow's stable descending (equal keys keep
 *                  their input order)
 *   idx_out        n uint32 row indices: idx_out[j] = input row of output row j
 *   sorted_out     optional (may be NULL): the key column in sorted order
 *   options        DS_SORT_* bits.  Unless DS_SORT_NO_PASS_SKIP is set the call
 *                  does ONE small host sync to skip radix passes whose digit is
 *                  constant over the whole column.  Unless DS_SORT_NO_PREFIX_REPAIR is
 *                  set it may also sort by the most significant digits only when the
 *                  histograms show that they separate almost all rows, then order the
 *                  remaining short runs of equal prefixes by the full key (verified on
 *                  device; a second 4-byte host sync; falls back to all passes if a run
 *                  is long).  Unless DS_SORT_NO_HYBRID is set, keys whose two leading bytes
 *                  spread evenly (and 26 M <= n <= 108 M) run only those two digits as global
 *                  passes and finish with one thread block per 16-bit bucket in shared
 *                  memory (same verification, same fallback).  Keys that fill only a
 *                  sub-range of the key space evenly (one rank's share of a range partition,
 *                  clustered ids) take the same plan with the bucket digits computed from the
 *                  key's position between the sampled minimum and maximum; every key is
 *                  checked against that range on the device and an outlier falls back.  The
 *                  result is identical in every mode.
 * float64 order: -inf < ... < -0.0 == 0.0 ... < +inf < NaN (NaNs last in ascending
 * and in descending order, like Arrow); -0.0 and 0.0 compare equal (stable). */
ds_status ds_sort_indices_i64(const int64_t* col, uint64_t n, int32_t descending,
                              uint32_t* idx_out, int64_t* sorted_out, int32_t options,
                              void* ws, size_t* ws_bytes, void* stream);
ds_status ds_sort_indices_u64(const uint64_t* col, uint64_t n, int32_t descending,
                              uint32_t* idx_out, uint64_t* sorted_out, int32_t options,
                              void* ws, size_t* ws_bytes, void* stream);
ds_status ds_sort_indices_f64(const double* col, uint64_t n, int32_t descending,
                              uint32_t* idx_out, double* sorted_out, int32_t options,
                              void* ws, size_t* ws_bytes, void* stream);

/* Stable argsort of a fixed_size_binary(width) key column (memcmp order), width in
 * 1..DS_MAX_KEY_BYTES -- the 10-byte GraySort key.  Same call sites as above.
 *   col          n*width bytes, rows packed (Arrow fixed_size_binary data buffer)
 *   sorted_out   optional: n*width bytes, the key column in sorted order */
ds_status ds_sort_indices_fsb(const uint8_t* col, uint32_t width, uint64_t n,
                              int32_t descending, uint32_t* idx_out, uint8_t* sorted_out,
                              int32_t options, void* ws, size_t* ws_bytes, void* stream);

/* Row gather dst[j] = src[idx[j]] for rows of row_bytes bytes.
 * Replaces Table.take / take_table at transform_pyarrow.py:16,46-67 and
 * ArrowBlockAccessor.take (_internal/arrow_block.py:258-267); one call per column.
 *   src_rows   rows in the source column (the kernels read whole aligned 16-byte
 *              vectors and need the end of the buffer to stay inside it)
 *   idx        n uint32 source row numbers, each < src_rows (not checked) */
ds_status ds_gather_rows(const void* src, uint32_t row_bytes, uint64_t src_rows,
                         const uint32_t* idx, uint64_t n, void* dst, void* stream);

/* The same gather when the source column lies in several buffers -- the blocks of a dataset --
 * so that sorting a dataset of many blocks does not first copy them into one
 * (pa.concat_tables at transform_pyarrow.py:70-75 / _internal/arrow_block.py:473-486 is free for
 * Arrow's chunked tables; a device column has to be contiguous or gathered like this).
 * idx[j] counts rows over the concatenation of the sources in the given order.
 *   srcs, src_rows   HOST arrays of nsrc device pointers / row counts (nsrc <= 64)
 * Supported for even row_bytes in 18..1024 that are not a multiple of 16, sources 2-byte and dst
 * 16-byte aligned (the asynchronous-copy kernel); otherwise DS_ERR_UNSUPPORTED -- concatenate. */
ds_status ds_gather_rows_multi(const void* const* srcs, const uint64_t* src_rows, uint32_t nsrc,
                               uint32_t row_bytes, const uint32_t* idx, uint64_t n, void* dst,
                               void* stream);

/* Partition scatter: copy the rows of a column, visited in partitioned order, to where
 * each partition is consumed.  Row j (in the order of `perm`, as written by
 * ds_partition_hist_scatter_*) belongs to partition d with part_offsets[d] <= j <
 * part_offsets[d+1] and is stored at dst_bases[d] + (j - part_offsets[d]) * row_bytes.
 * dst_bases[d] may be peer-mapped memory of another GPU (NVLink) or local memory, so one
 * launch performs the map-side slicing AND the transfer of the all-to-all exchange.
 * Replaces the M x N map-output objects and their pull by the reducers
 * (_internal/shuffle.py:81-112; src/ray/object_manager/object_manager.cc:336-559).
 *   part_offsets   HOST array of nparts+1 row offsets, part_offsets[nparts] == n
 *   dst_bases      HOST array of nparts device pointers (nparts <= 16) */
ds_status ds_scatter_rows_to_parts(const void* src, uint32_t row_bytes, uint64_t src_rows,
                                   const uint32_t* perm, uint64_t n, uint32_t nparts,
                                   const uint64_t* part_offsets, void* const* dst_bases,
                                   void* stream);

/* Streaming split -- the map side of the exchange as ONE pass over the input, rows kept in
 * input order (no permutation, no partitioned copy of the block).
 * Replaces the per-mapper slicing of sort_and_partition (_internal/arrow_block.py:387-393),
 * the M x N map outputs of SimpleShufflePlan.execute (_internal/shuffle.py:81-104) and their
 * pull by the reduce tasks (_internal/shuffle.py:106-112;
 * src/ray/object_manager/object_manager.cc:336-559).
 *
 * ds_split_plan: destination of every row and the per-tile placement table.
 *   key_kind          DS_KEY_*; width = key bytes for DS_KEY_FSB (ignored otherwise)
 *   valid             optional validity bytes (NULL = no nulls); null keys go to the LAST
 *                     destination (the reducer's sort puts them at the end)
 *   boundaries        nb keys (device) in the column's own encoding, ordered like the sort
 *                     (same rule as ds_partition_hist_scatter_*); nb + 1 <= DS_MAX_SPLIT_PARTS
 *   dest_out          n bytes: destination of row i
 *   tile_offsets_out  capacity ds_split_table_words(n, nb+1) uint32.  The first
 *                     (nb+1) * (ntiles+1) words, ntiles = ceil(n / DS_SPLIT_TILE_ROWS), are the
 *                     table: entry [d*(ntiles+1) + t] = rows of destination d in tiles < t;
 *                     entry [d*(ntiles+1) + ntiles] = all rows of destination d.  The rest is
 *                     scratch of the scan.
 *   counts_out        nb+1 uint64 rows per destination (device)
 * ds_split_rows: moves one column.  Row i (destination d = dest[i]) is stored at
 *   dst_bases[d] + (number of earlier rows with destination d) * row_bytes; dst_bases[d] may
 *   be peer-mapped memory of another GPU (the rows then cross NVLink as bulk stores issued
 *   from shared memory) or local memory.
 *   dst_bases   HOST array of nparts device pointers
 *   dest_mask   bit d set: the rows of destination d are moved by this call; rows of the other
 *               destinations are left where they are (0 = every destination).  A plan that
 *               sends its key ranges in rounds calls once per round with that round's mask.
 *   max_ctas    upper bound on the persistent grid (0 = as many CTAs per SM as fit, up to 3),
 *               so that a caller can leave room for work that runs beside the transfer
 *   row_bytes <= DS_SPLIT_MAX_ROW_BYTES. */
#define DS_KEY_I64 0
#define DS_KEY_U64 1
#define DS_KEY_F64 2
#define DS_KEY_FSB 3
#define DS_SPLIT_TILE_ROWS 256
#define DS_MAX_SPLIT_PARTS 32
#define DS_SPLIT_MAX_ROW_BYTES 200
uint64_t ds_split_table_words(uint64_t n, uint32_t nparts);
ds_status ds_split_plan(int32_t key_kind, const void* col, uint32_t width, const uint8_t* valid,
                        uint64_t n, const void* boundaries, uint32_t nb, int32_t descending,
                        uint8_t* dest_out, uint32_t* tile_offsets_out, uint64_t* counts_out,
                        void* stream);
ds_status ds_split_rows(const void* src, uint32_t row_bytes, uint64_t n, const uint8_t* dest,
                        const uint32_t* tile_offsets, uint32_t nparts, void* const* dst_bases,
                        uint32_t dest_mask, uint32_t max_ctas, void* stream);

/* ----------------------------------------------------------- partition ---- */

/* Cut positions of a SORTED key column at the given boundaries.
 * Replaces np.searchsorted(table[col], boundaries) (ascending, side="left") and
 * n - np.searchsorted(..., sorter=arange(n-1,-1,-1)) (descending) at
 * _internal/arrow_block.py:380-386.
 *   boundaries   nb keys in the same encoding as the column (ascending order for
 *                ascending sorts, descending order for descending sorts)
 *   bounds_out   nb uint64 row offsets (device) */
ds_status ds_searchsorted_i64(const int64_t* sorted_col, uint64_t n, const int64_t* boundaries,
                              uint32_t nb, int32_t descending, uint64_t* bounds_out, void* stream);
ds_status ds_searchsorted_f64(const double* sorted_col, uint64_t n, const double* boundaries,
                              uint32_t nb, int32_t descending, uint64_t* bounds_out, void* stream);
ds_status ds_searchsorted_fsb(const uint8_t* sorted_col, uint32_t width, uint64_t n,
                              const uint8_t* boundaries, uint32_t nb, int32_t descending,
                              uint64_t* bounds_out, void* stream);

/* Stable range partition of an UNSORTED key column: the map side of the exchange.
 * Same partition rule as sort_and_partition (_internal/arrow_block.py:373-394):
 * ascending  bucket(x) = #{i : boundaries[i] <= x};
 * descending bucket(x) = #{i : boundaries[i] >  x}.
 *   counts_out   nb+1 uint64 rows per bucket (device)
 *   offsets_out  nb+1 uint64 exclusive prefix of counts_out (device)
 *   perm_out     n uint32: rows grouped by bucket, input order kept inside a bucket
 * Replaces the per-mapper slicing at _internal/arrow_block.py:387-393 when the plan
 * partitions before sorting (multi-GPU plan). */
ds_status ds_partition_hist_scatter_i64(const int64_t* col, uint64_t n, const int64_t* boundaries,
                                        uint32_t nb, int32_t descending, uint64_t* counts_out,
                                        uint64_t* offsets_out, uint32_t* perm_out, void* ws,
                                        size_t* ws_bytes, void* stream);
ds_status ds_partition_hist_scatter_fsb(const uint8_t* col, uint32_t width, uint64_t n,
                                        const uint8_t* boundaries, uint32_t nb, int32_t descending,
                                        uint64_t* counts_out, uint64_t* offsets_out,
                                        uint32_t* perm_out, void* ws, size_t* ws_bytes,
                                        void* stream);

/* ------------------------------------------------------------- sampling ---- */

/* k sample rows of a column, without replacement: sample j is taken from stratum
 * [floor(j*n/k), floor((j+1)*n/k)) at a Philox4x32-10(seed, j) offset.
 * Replaces random.sample(range(n), k) + take at _internal/arrow_block.py:269-272
 * (reference sampling is unseeded; which rows are drawn is not part of its contract).
 * Requires 1 <= k <= n.  out: k*row_bytes bytes. */
ds_status ds_sample_rows(const void* col, uint32_t row_bytes, uint64_t n, uint32_t k,
                         uint64_t seed, void* out, void* stream);

/* -------------------------------------------------------- random_shuffle ---- */

/* A seeded uniform permutation of range(n): Philox4x32-10(seed, subsequence, row)
 * 64-bit keys, stable radix sort, row index as payload.
 * Replaces np.random.RandomState(seed).permutation(n) at
 * _internal/arrow_block.py:184-186 (and the reducer-side call,
 * _internal/shuffle_and_partition.py:86-89).  NOT bit-equal to MT19937: the
 * contract kept is "a permutation, deterministic per (seed, subsequence, n)". */
ds_status ds_random_permutation(uint64_t seed, uint64_t subsequence, uint64_t n,
                                uint32_t* perm_out, void* ws, size_t* ws_bytes, void* stream);

/* --------------------------------------------------------------- groupby ---- */

/* Run starts of a SORTED key column: seg_starts_out[s] = first row of the s-th run
 * of equal keys, *nseg_out = number of runs.  Replaces the per-row group scan of
 * ArrowBlockAccessor.combine (_internal/arrow_block.py:417-442) and the key-change
 * test of aggregate_combined_blocks (_internal/arrow_block.py:520-547).
 *   seg_starts_out  capacity n+1 uint32; entry nseg is set to n
 *   nseg_out        one uint32 (device) */
ds_status ds_segment_heads_i64(const int64_t* sorted_keys, uint64_t n, uint32_t* seg_starts_out,
                               uint32_t* nseg_out, void* ws, size_t* ws_bytes, void* stream);

/* Same result as ds_segment_heads_i64, for callers that cannot hand over a workspace
 * because they cannot allocate device memory at the point of the call: a reduce task
 * running inside a host runtime whose HBM is fully owned by its own pool, a binding
 * (ctypes / cgo / JNI) that has the input and output buffers but no allocator handle,
 * or a worker that has planned its HBM to the last block and must not see an
 * unplanned cudaMalloc fail or stall the device.  This entry point has no counterpart
 * in the reference.
 *   - scratch (one counter per 2048 keys) is a fixed array inside the library image,
 *     so n is limited to DS_SEGMENT_HEADS_STATIC_MAX_ROWS; larger n is rejected with
 *     DS_ERR_TOO_LARGE.  Each GPU that loads the library holds the array (~256 KiB)
 *     until the process exits.
 *   - the call makes no cudaMalloc / cudaFree and calls no allocating routine.
 *   - the first call on a device looks up the array's device address (a one-time
 *     set-up step with that device current); make that first call before any stream
 *     capture begins.
 *     The address is cached per device for the life of the process: after a
 *     cudaDeviceReset the cached value is stale and the entry point must not be used
 *     again in that process.
 *   - all calls on a device share the one array and find it as the previous call
 *     left it; nothing in it needs clearing (every counter is rewritten before it is
 *     read).  Two calls whose kernels overlap on the GPU (different streams, no
 *     ordering between them) race on it: serialise them.
 *   seg_starts_out  capacity n+1 uint32; entry nseg is set to n
 *   nseg_out        one uint32 (device) */
#define DS_SEGMENT_HEADS_STATIC_MAX_ROWS (1u << 27)
ds_status ds_segment_heads_i64_static(const int64_t* sorted_keys, uint64_t n,
                                      uint32_t* seg_starts_out, uint32_t* nseg_out,
                                      void* stream);

/* Per-run reduction of a value column that is still in INPUT order: run s reduces
 * vals[idx[i]] for i in [seg_starts[s], seg_starts[s+1]).  idx may be NULL (values
 * already in sorted order).  Replaces pyarrow.compute.sum/min/max on a group slice
 * (_internal/arrow_block.py:311-329 via aggregate.py:97-211) and the accumulator
 * merges a1+a2 of aggregate_combined_blocks (aggregate.py:107,119,185-187).
 *   valid       optional (may be NULL): one byte per INPUT row, 0 = null.  Null rows
 *               contribute the op's identity, i.e. they are skipped
 *               (skip_nulls=True at _internal/arrow_block.py:309); use
 *               ds_segmented_count_valid to learn how many rows of a run were non-null
 *   seg_starts  nseg+1 entries as written by ds_segment_heads_i64
 *   nseg        HOST value (the caller reads *nseg_out back to size `out`)
 *   op          DS_OP_SUM / DS_OP_MIN / DS_OP_MAX / DS_OP_ABSMAX
 * Summation order inside a run is fixed by (n, seg_starts) alone: results are
 * reproducible run to run (no floating-point atomics). */
ds_status ds_segmented_reduce_f64(const double* vals, const uint8_t* valid, const uint32_t* idx,
                                  const uint32_t* seg_starts, uint32_t nseg, uint64_t n,
                                  int32_t op, double* out, void* ws, size_t* ws_bytes,
                                  void* stream);
ds_status ds_segmented_reduce_i64(const int64_t* vals, const uint8_t* valid, const uint32_t* idx,
                                  const uint32_t* seg_starts, uint32_t nseg, uint64_t n,
                                  int32_t op, int64_t* out, void* ws, size_t* ws_bytes,
                                  void* stream);

/* Non-null rows per run: out[s] = #{i in run s : valid[idx[i]] != 0} (all rows if valid is
 * NULL).  Replaces ArrowBlockAccessor.count (_internal/arrow_block.py:274-288, used by
 * Mean at aggregate.py:191) and the has_data bookkeeping of
 * _internal/null_aggregate.py:191-225. */
ds_status ds_segmented_count_valid(const uint8_t* valid, const uint32_t* idx,
                                   const uint32_t* seg_starts, uint32_t nseg, uint64_t n,
                                   int64_t* out, void* ws, size_t* ws_bytes, void* stream);

/* Row counts per run: out[s] = seg_starts[s+1] - seg_starts[s].
 * Replaces Count.accumulate_block (aggregate.py:97-109). */
ds_status ds_segment_counts(const uint32_t* seg_starts, uint32_t nseg, int64_t* out,
                            void* stream);

/* Per-run sum of addend[j] + weights[j] * (vals[j] - means[s])^2 over the non-null rows of
 * run s (weights and addend optional).  With neither it is the block-level half of Std,
 * BlockAccessor.sum_of_squared_diffs_from_mean (_internal/arrow_block.py:331-351, used at
 * aggregate.py:262); with weights = counts and addend = M2 over partial rows (vals = partial
 * means) it is the merge of Std accumulators (aggregate.py:236-249), evaluated against the
 * merged mean instead of pairwise. */
ds_status ds_segmented_sumsqdiff_f64(const double* vals, const uint8_t* valid, const int64_t* weights,
                                     const double* addend, const uint32_t* idx,
                                     const uint32_t* seg_starts, uint32_t nseg, uint64_t n,
                                     const double* means, double* out, void* ws, size_t* ws_bytes,
                                     void* stream);
ds_status ds_segmented_sumsqdiff_i64(const int64_t* vals, const uint8_t* valid, const uint32_t* idx,
                                     const uint32_t* seg_starts, uint32_t nseg, uint64_t n,
                                     const double* means, double* out, void* ws, size_t* ws_bytes,
                                     void* stream);

/* out[i] = a[i] * w[i]: partial mean x partial count = partial sum (aggregate.py:246). */
ds_status ds_scale_f64_by_i64(const double* a, const int64_t* w, uint64_t n, double* out,
                              void* stream);

/* Std finalize (aggregate.py:270-276): 0.0 if count < 2 else sqrt(m2 / (count - ddof)). */
ds_status ds_finalize_std(const double* m2, const int64_t* count, uint64_t n, int32_t ddof,
                          double* out, void* stream);

/* Validity of a finalized aggregate per group (null policy of
 * _internal/null_aggregate.py:12-26, _null_wrap_finalize :228-251): with ignore_nulls the
 * result is a value unless every row of the group was null; without, only if none was.
 *   nvalid / nrows   non-null rows and all rows per group (int64)
 *   out              one byte per group, 1 = value, 0 = null */
ds_status ds_finalize_validity(const int64_t* nvalid, const int64_t* nrows, uint64_t n,
                               int32_t ignore_nulls, uint8_t* out, void* stream);

/* Mean finalize out[s] = sum[s] / count[s] (aggregate.py:209).  sum_is_i64 selects
 * the accumulator type of `sum` (int64 sums of int64 columns stay integral until
 * the true division, like Python's int / int). */
ds_status ds_finalize_mean(const void* sum, int32_t sum_is_i64, const int64_t* count,
                           uint64_t n, double* out, void* stream);

/* ------------------------------------------------------- GraySort records ---- */

/* On-disk GraySort layout (records of record_bytes back to back, key first;
 * experimental/raysort/constants.py:8 RECORD_SIZE = 100) <-> the two Arrow columns
 * {key: fixed_size_binary(key_bytes), payload: fixed_size_binary(record_bytes-key_bytes)}.
 * The reference reads and writes such files through gensort/valsort subprocesses
 * (experimental/raysort/main.py:148-176,399-432); there is no in-tree parser to replace. */
ds_status ds_split_records(const uint8_t* records, uint32_t record_bytes, uint32_t key_bytes,
                           uint64_t n, uint8_t* key_out, uint8_t* payload_out, void* stream);
ds_status ds_join_records(const uint8_t* key, const uint8_t* payload, uint32_t record_bytes,
                          uint32_t key_bytes, uint64_t n, uint8_t* records_out, void* stream);

/* valsort-style summary of a (key, payload) block, ADDED to out3 (device, zero it first):
 * out3[0] += sum over records of CRC-32(record) (order independent, so equal before and
 * after a correct sort), out3[1] += records whose key is below the previous record's,
 * out3[2] += records whose key equals the previous record's. */
ds_status ds_validate_records(const uint8_t* key, const uint8_t* payload, uint32_t record_bytes,
                              uint32_t key_bytes, uint64_t n, uint64_t* out3, void* stream);

/* ----------------------------------------- variable-length keys and columns ---- */

/* Arrow `binary` / `string` (and large_*) columns on the path: n+1 offsets into a byte
 * buffer.  On the device offsets are int64 whatever the Arrow offset width and may start
 * anywhere in the buffer (a slice keeps its parent's bytes).  In the reference these
 * columns go through pyarrow on the CPU: `take` for every permutation of rows
 * (_internal/arrow_block.py:258-267, arrow_ops/transform_pyarrow.py:46-60),
 * `pyarrow.compute.sort_indices` for the sort (transform_pyarrow.py:15,74 called from
 * arrow_block.py:353-394,474-486) and `np.searchsorted` on the sorted key column for the
 * partition cuts (arrow_block.py:376-386).
 *
 * take = two calls: ds_varlen_take_offsets writes the n+1 offsets of the rows idx[0..n)
 * (idx NULL = rows 0..n-1 in order, i.e. a compaction of a slice) by a 64-bit scan of their
 * lengths (workspace protocol); the caller reads dst_offsets[n], allocates that many bytes
 * and calls ds_varlen_take_bytes. */
ds_status ds_varlen_take_offsets(const int64_t* src_offsets, uint64_t src_rows, const uint32_t* idx,
                                 uint64_t n, int64_t* dst_offsets, void* ws, size_t* ws_bytes,
                                 void* stream);
ds_status ds_varlen_take_bytes(const int64_t* src_offsets, const uint8_t* src_data,
                               const uint32_t* idx, uint64_t n, const int64_t* dst_offsets,
                               uint8_t* dst_data, void* stream);

/* The 64-bit parts a variable-length key is radix sorted by (least significant first: the
 * length, then 8-byte chunks from the last to the first; every part sorted with the stable
 * ds_sort_indices_u64 in the direction of the sort).  Together they order rows like
 * memcmp-then-length, which is pyarrow's sort_indices order for binary and string.
 *   idx             rows in their current order (NULL = identity); part i is taken from
 *                   row idx[i], so each round's keys are already permuted
 *   lengths_out[i]  byte length of row idx[i]; *max_len_out (device) = longest row
 *   keys_out[i]     bytes [byte_offset, byte_offset+8) of row idx[i], first byte most
 *                   significant, zero padded past the end of the row */
ds_status ds_varlen_key_lengths(const int64_t* offsets, const uint32_t* idx, uint64_t n,
                                uint64_t* lengths_out, uint64_t* max_len_out, void* stream);
ds_status ds_varlen_key_chunk(const int64_t* offsets, const uint8_t* data, const uint32_t* idx,
                              uint64_t n, uint64_t byte_offset, uint64_t* keys_out, void* stream);

/* np.searchsorted(sorted_keys, boundaries, side="left") for a sorted variable-length column
 * (arrow_block.py:376-386; descending counts the rows greater than each boundary, which is
 * the reference's `num_rows - searchsorted(reversed, side="right")`).  Boundaries are given
 * in the same layout (nb+1 offsets + bytes, on the device). */
ds_status ds_searchsorted_varlen(const int64_t* sorted_offsets, const uint8_t* sorted_data,
                                 uint64_t n, const int64_t* bound_offsets,
                                 const uint8_t* bound_data, uint32_t nb, int32_t descending,
                                 int64_t* out, void* stream);

/* Runs of equal keys of a key-SORTED column as an int64 column: group_ids_out[i] = number of
 * key changes in rows (0, i].  It has the same runs as the key, so ds_segment_heads_i64 and
 * the segmented reductions work on it for every key type; this is what replaces the
 * row-by-row `next_row[key] != current key` scan of ArrowBlockAccessor.combine
 * (arrow_block.py:417-440) for keys that are not int64.
 *   ds_group_ids_fixed   rows of `width` bytes compared bytewise, or (is_float64) doubles
 *                        compared numerically (-0.0 == 0.0; a NaN equals nothing, as in
 *                        Python)
 *   ds_group_ids_varlen  rows compared by length and bytes */
ds_status ds_group_ids_fixed(const void* sorted_col, uint32_t width, int32_t is_float64, uint64_t n,
                             int64_t* group_ids_out, void* ws, size_t* ws_bytes, void* stream);
ds_status ds_group_ids_varlen(const int64_t* sorted_offsets, const uint8_t* sorted_data, uint64_t n,
                              int64_t* group_ids_out, void* ws, size_t* ws_bytes, void* stream);

/* ------------------------------------------------------- narrow numeric keys ---- */

/* pyarrow's sort_indices / searchsorted take a key column of any numeric type
 * (transform_pyarrow.py:15, arrow_block.py:380-386); the kernels above take int64, uint64,
 * float64 and short fixed_size_binary.  ds_widen_keys converts a narrower numeric column into
 * the 64-bit type that orders and ties identically (integers -> int64, uint64 -> int64 biased
 * by 2^63, float16/float32 -> float64, NaNs stay NaNs), which then stands in for the key in the calls above while the
 * original column is what the row gather moves.  out64: n 8-byte values. */
#define DS_KEY_INT8 0
#define DS_KEY_INT16 1
#define DS_KEY_INT32 2
#define DS_KEY_UINT8 3
#define DS_KEY_UINT16 4
#define DS_KEY_UINT32 5
#define DS_KEY_FLOAT32 6
#define DS_KEY_FLOAT16 7
#define DS_KEY_UINT64 8 /* -> int64 with the top bit flipped (same order) */
ds_status ds_widen_keys(const void* col, int32_t src_type, uint64_t n, void* out64, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DS_SHUFFLE_H_ */
