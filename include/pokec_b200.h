/* pokec_b200.h -- C ABI of libpokec_b200.so: the B200 (sm_100a) scoring path behind the reference's
 * C++ `Recommender` (pymlex/recommendation-system-pokec, include/recommender.h:17-71).
 *
 * Everything the reference's Recommender computes on the hot path -- Fill-Aware Similarity (FAS),
 * per-column idf, candidate scoring with top-k, collaborative friend / club propagation -- is
 * reached through the entry points below.  Plain pointers and sizes only; all array arguments are
 * HOST pointers unless the name starts with `d_` (device pointers, used by the multi-GPU plumbing).
 * The library copies what it keeps.  There is NO CPU fallback: every scoring entry point needs a
 * CUDA device and fails with PK_ERR_CUDA otherwise.
 *
 * Threading: one pk_store may be used from one host thread at a time (the reference's callers are
 * single-threaded: src/api_cli.cpp:176, python/app.py:56-59).  Nothing is ever written to stdout
 * (stdout is api_cli's JSON channel, src/api_cli.cpp:172-255); diagnostics go to pk_last_error().
 *
 * Reference-side binding: see INTEGRATION.md (replacement translation units for recommender*.cpp
 * that implement the unchanged include/recommender.h on top of this ABI).
 */
#ifndef POKEC_B200_H
#define POKEC_B200_H

#include <stdint.h>
#include "pokec_world.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pk_store pk_store;

enum {
    PK_OK = 0,
    PK_ERR_ARG = -1,         /* bad argument (null pointer, negative size, ...) */
    PK_ERR_CUDA = -2,        /* CUDA runtime failure or no device */
    PK_ERR_UNSUPPORTED = -3, /* outside the supported envelope (T > 64, token id >= 2^26, full-scan k > 256, list-mode k > 8192) */
    PK_ERR_STATE = -4        /* call order (no world loaded, ...) */
};

/* ---- life cycle ------------------------------------------------------------------------------ */

/* Message of the last failing call on this thread (never NULL). */
const char* pk_last_error(void);

/* ABI version of this header (PK_ABI_VERSION). */
#define PK_ABI_VERSION 1
int pk_abi_version(void);

/* Number of CUDA devices visible (0 when there is none; never fails). */
int pk_device_count(void);

/* Replaces `Recommender::Recommender(profiles*, adj_list*)` (src/recommender.cpp:9-16): an empty
 * store bound to CUDA device `device`. */
int pk_store_create(int device, pk_store** out);
void pk_store_destroy(pk_store* s);

/* Uploads the flat form of `profiles` + `adj_list` (include/user_profile.h:10-20,
 * include/recommender.h:46-49) and the normalizers carried by `w` (may be all-absent).  Token ids
 * must be < 2^26 and n_text_cols <= 64.  Cells need not be sorted (sorted on ingest). */
int pk_store_load_world(pk_store* s, const pk_world* w);

/* Candidate shard for the full scan: rows [row_begin,row_end) of the loaded world (world order).
 * Queries may be any loaded user.  Default: all rows.  (north_star: candidates partitioned by
 * user-id range across the GPUs of one box.) */
int pk_store_set_shard(pk_store* s, int64_t row_begin, int64_t row_end);

/* Tuning knobs of a store.
 *   PK_OPT_PREFILTER_RESERVE_SMS  the persistent pre-filter of the full scan normally takes every SM for the length of the
 *                                 scan; a caller that keeps two batches in flight on two stores / streams can ask it to
 *                                 leave `value` SMs free, so that the previous batch's merges, re-score and NCCL
 *                                 all-gathers overlap the scan instead of queueing behind it (0 = default). */
enum { PK_OPT_PREFILTER_RESERVE_SMS = 1 };
int pk_store_set_option(pk_store* s, int which, int64_t value);

/* Replaces set_field_normalizers + set_column_normalizers (src/recommender.cpp:27-33): `n_fields`
 * = 7 + T entries, the 7 fixed keys (public, gender, completion, age, region, clubs, friends)
 * followed by the text columns.  has[i]==0 means "no entry in the map" (z = 6(s-0.5),
 * src/recommender_similarity.cpp:28-36). */
int pk_set_normalizers(pk_store* s, const float* mean, const float* sd, const uint8_t* has, int n_fields);

/* Replaces Recommender::compute_idf_from_profiles (src/recommender.cpp:43-66): per column document
 * frequencies on the device, idf = logf(1 + N/(1+df)) (float, evaluated with the host libm so the
 * table is bit-identical to the reference's).  double_log != 0 selects TFIDFIndex::build's variant
 * (src/tfidf_index.cpp:7-30: (float)log(1.0 + N/(1.0+df)) in double). */
int pk_compute_idf(pk_store* s, int double_log);

/* Replaces Recommender::set_tfidf_index for one column (src/recommender.cpp:35-37): the column's
 * idf map becomes exactly {token[i] -> idf[i]}.  Tokens absent from the map score with idf 1.0f
 * (src/recommender.cpp:78).  n < 0 erases the column's entry: that column then uses the raw-count
 * cosine (src/recommender.cpp:141-163 via src/recommender_similarity.cpp:99-104). */
int pk_set_idf(pk_store* s, int col, const int32_t* token, const float* idf, int64_t n);

/* Reads back a column's idf map (the public member idf_per_col, include/recommender.h:53), sorted
 * by token id.  pk_idf_size returns -1 in *n when the column has no entry.  df may be NULL; it is 0
 * for entries that came from pk_set_idf. */
int pk_idf_size(pk_store* s, int col, int64_t* n);
int pk_get_idf(pk_store* s, int col, int32_t* token, float* idf, int32_t* df, int64_t cap);

/* Replaces one row of the caller-owned adjacency after the caller mutated it behind the
 * Recommender (src/test.cpp:59-73).  n == 0 keeps an empty row; the row is created if missing. */
int pk_set_adj_row(pk_store* s, int32_t uid, const int32_t* nb, int64_t n);

/* ---- scoring --------------------------------------------------------------------------------- */

/* Replaces Recommender::profile_similarity (src/recommender_similarity.cpp:10-128) for n pairs
 * (A = a_uid[i], B = b_uid[i]).  out[i] = -1 when either uid has no profile. */
int pk_pair_similarity(pk_store* s, int64_t n, const int32_t* a_uid, const int32_t* b_uid, float* out);

/* Per-field float similarities of n pairs: out[i*(3+T) + j], j = 0 region, 1 clubs, 2 friends,
 * 3+t text column t; -1 = field absent on either side.  (The float rounding points of
 * src/recommender.cpp:116,127,138.) */
int pk_pair_fields(pk_store* s, int64_t n, const int32_t* a_uid, const int32_t* b_uid, float* out);

/* Replaces compute_column_normalizers (src/utils.cpp:155-240) for a caller-supplied sample of n user pairs (the
 * reference draws them with mt19937(12345) from its unordered_map's iteration order, which only the caller's C++
 * side can reproduce): per field -- the 7 fixed keys then the text columns, as in pk_set_normalizers -- the mean and
 * the sample standard deviation (1.0 when 0 or n < 2) of the RAW similarities, 0 where a field is absent on a side,
 * raw-count cosine (src/utils.cpp:99-121) for every text column whatever idf tables are set.  The per-pair
 * similarities come from the device; the sums run in pair order in double, so the floats are bit-identical to the
 * reference's.  Fails with PK_ERR_ARG when a pair names a user without a profile (the reference's at() throws). */
int pk_normalizers_from_pairs(pk_store* s, int64_t n, const int32_t* a_uid, const int32_t* b_uid, float* out_mean,
                              float* out_sd);

/* Set-overlap counts of n pairs (which: 0 clubs, 1 friends), the integer numerator of
 * vec_set_similarity (src/recommender.cpp:119-128); -1 when either uid has no profile. */
int pk_pair_overlap(pk_store* s, int64_t n, const int32_t* a_uid, const int32_t* b_uid, int which, int32_t* out);

/* The north-star path: FAS of every query against EVERY candidate profile of the shard, top-k by
 * the reference comparator (score desc, uid asc; src/recommender_graph.cpp:97-100).
 * exclude_mode: PK_EXCLUDE_* (pokec_world.h).  out_id/out_score are [nq][k] (unused slots: -1 / 0),
 * out_cnt[nq] the number of valid entries.  Unknown query uid -> cnt 0. */
int pk_fullscan_topk(pk_store* s, int nq, const int32_t* q_uid, int k, int exclude_mode,
                     int32_t* out_id, float* out_score, int32_t* out_cnt);

/* The same call in three stages, for callers that pipeline batches or time the device work alone:
 *   prepare  host planning (queries -> groups of 32 -> work items) and upload of the plan
 *   execute  query-table build, scan and merge kernels; asynchronous on the store's stream; may be
 *            repeated (it recomputes everything from device-resident inputs)
 *   fetch    synchronise, copy the merged keys to the host and decode them
 *   keys     asynchronous device-to-device copy of the merged keys ([nq][k] uint64) into d_keys */
int pk_fullscan_prepare(pk_store* s, int nq, const int32_t* q_uid, int k, int exclude_mode);
int pk_fullscan_execute(pk_store* s);
int pk_fullscan_fetch(pk_store* s, int32_t* out_id, float* out_score, int32_t* out_cnt);
int pk_fullscan_keys(pk_store* s, uint64_t* d_keys);

/* Batched form of the reference's recommend calls (include/recommender.h:24-28):
 *   PK_MODE_GRAPH  recommend_graph_registration == recommend_by_interest (src/recommender_graph.cpp:33-103,224-227)
 *   PK_MODE_COLLAB recommend_collaborative (src/recommender_graph.cpp:105-222)
 *   PK_MODE_CLUBS  recommend_clubs_collab (src/recommender_clubs.cpp:10-73; out_id = club ids, limit ignored)
 * Same output layout as pk_fullscan_topk. */
int pk_recommend(pk_store* s, int mode, int nq, const int32_t* q_uid, int k, int limit,
                 int32_t* out_id, float* out_score, int32_t* out_cnt);

/* Per-edge FAS cache.  Both weights of recommend_clubs_collab are FAS values on adjacency
 * entries -- w_f = FAS(u, f), f in adj[u] (src/recommender_clubs.cpp:23-28), and s = FAS(f, fof), fof in adj[f]
 * (:56) -- and do not depend on the query.  pk_edge_cache_build scores every loaded adjacency entry once with the
 * exact kernel (-1 where a side has no profile) and keeps the floats on the device; pk_recommend(PK_MODE_CLUBS)
 * then only enumerates edges and scatters (same visiting order, same floats, bit-identical results).  The cache is
 * dropped by pk_store_load_world, pk_set_normalizers, pk_compute_idf and pk_set_idf, and bypassed while rows
 * replaced with pk_set_adj_row exist.  pk_edge_cache_row reads the cached values of uid's adjacency row (in row
 * order; *n = row length). */
int pk_edge_cache_build(pk_store* s);
int pk_edge_cache_drop(pk_store* s);
int pk_edge_cache_row(pk_store* s, int32_t uid, float* out, int64_t cap, int64_t* n);

/* Candidate id lists of one user (integer contract): PK_MODE_GRAPH = gather_candidates_local
 * (src/recommender_graph.cpp:10-31), PK_MODE_COLLAB = the FoF expansion of :114-125.
 * Returns the list length in *n (may exceed cap; only cap entries are written). */
int pk_candidates(pk_store* s, int mode, int32_t uid, int limit, int32_t* out, int64_t cap, int64_t* n);

/* The same lists for nq users in one call, expanded ON THE DEVICE (adjacency CSR resident in HBM, one warp per user:
 * ordered de-duplication against a per-warp hash set, truncation at `limit`) -- the enumeration pk_recommend uses for
 * batches of PK_MODE_GRAPH / PK_MODE_COLLAB.  out_ids is [nq][limit] (only the first out_cnt[q] entries of a row are
 * written), same ids in the same order as pk_candidates.  PK_ERR_UNSUPPORTED when the device path cannot take the call:
 * limit outside [1, 16384], a neighbour row longer than 32768, or rows replaced with pk_set_adj_row (the device holds
 * the loaded adjacency; single-user drop-in calls keep walking the caller's live rows on the host). */
int pk_candidates_batch(pk_store* s, int mode, int nq, const int32_t* q_uid, int limit, int32_t* out_ids, int32_t* out_cnt);

/* pk_candidates_batch on DEVICE buffers, allocating nothing.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same kernel, same lists; its scratch (per-warp hash sets, epochs, work-queue cursor, the
 * [nq] row output) lives in one file-scope static device array of about 4 MB that every GPU loading the library holds
 * until exit, which bounds the call: limit in [1, 992], nq <= 8192 (PK_ERR_UNSUPPORTED beyond, with a message).
 * d_q_uid [nq], d_out_ids [nq][limit] and d_out_cnt [nq] are DEVICE pointers on the store's device.  The adjacency must
 * already be resident (PK_ERR_STATE otherwise: call pk_candidates_batch or a batched pk_recommend once first).  No
 * cudaMalloc / cudaFree / allocating library call is made.  Asynchronous on the store's stream: the inputs must be
 * complete before the call, pk_sync() before the outputs are read.  Every call on a device shares the array: calls that
 * overlap on the GPU (two stores or streams of one device) race -- serialise them. */
int pk_candidates_batch_noalloc(pk_store* s, int mode, int nq, const int32_t* d_q_uid, int limit, int32_t* d_out_ids,
                                int32_t* d_out_cnt);

/* ---- multi-GPU plumbing (one process per GPU; the all-gather itself is the caller's NCCL) ----- */

/* Local top-k of the shard as packed sort keys on the DEVICE: d_keys is a device buffer of
 * nq*k uint64 ((float bits << 33) | (~(uid ^ 0x80000000) << 1) | 1; 0 = empty slot).  Asynchronous
 * on the store's stream: call pk_sync() before another stream or library reads d_keys. */
int pk_fullscan_local_keys(pk_store* s, int nq, const int32_t* q_uid, int k, int exclude_mode, uint64_t* d_keys);

/* Merges n_lists key lists ([n_lists][nq][k] on the device, e.g. the all-gathered per-rank lists)
 * into the final host result.  With out_id, out_score and out_cnt all NULL only the device merge is enqueued
 * (asynchronous on the store's stream; nothing is fetched) -- for callers that time or pipeline the device work. */
int pk_merge_keys(pk_store* s, const uint64_t* d_keys, int n_lists, int nq, int k,
                  int32_t* out_id, float* out_score, int32_t* out_cnt);

/* The two-stage full scan split at its exchange point, for N candidate shards (one rank each) that share the
 * exact re-score instead of each re-scoring its own shortlist:
 *   execute_shortlists  table build + fp32 pre-filter of the shard: the shard's approximate top-Ks per query
 *                       (without the pre-filter -- pk_stat(PK_STAT_SHORTLIST) == 0 -- the whole exact scan runs here)
 *   shortlist_keys      asynchronous copy of those [nq][Ks] keys (Ks = pk_stat(PK_STAT_SHORTLIST)) into d_short,
 *                       the buffer the caller all-gathers
 *   execute_finish      merges n_lists gathered shortlists ([n_lists][nq][Ks] on the device; NULL = the local one),
 *                       proves every query's shortlist, re-scores with the reference's fp64 arithmetic only the
 *                       query groups rank `rank` of `n_ranks` owns (every rank holds the whole packed store), and
 *                       runs the exact fallback of unproven queries over its own shard.  pk_fullscan_keys then
 *                       yields this rank's [nq][k] keys (empty lists for queries it does not own) for the final
 *                       all-gather + pk_merge_keys.
 * pk_fullscan_execute == execute_shortlists + execute_finish(NULL, 1, 0, 1).  All asynchronous on the store's stream. */
/* Pruning thresholds across ranks, over NVLink peer memory.  The pre-filter skips a candidate whose header-only upper
 * bound is below theta = (k-th best approximate score so far) - 2 margin for every query of a group.  Every rank's k-th
 * best is a lower bound of the merged k-th best, so another rank's theta is valid on this shard too: the kernel polls the
 * peers' threshold arrays while it scans (no collective, no extra launch) and accepts entries tagged with the same plan
 * fingerprint and execution count.
 *   theta_export   the cudaIpc handle (64 bytes) of this store's threshold array, to be all-gathered by the caller
 *   theta_peers    maps the other ranks' arrays: `handles` = n_ranks x 64 bytes in rank order (own entry ignored); or, for
 *                  stores of the same process, `d_ptrs` = their pk_fullscan_theta_ptr values (handles NULL).  Resets the
 *                  execution counter -- call it on every rank at the same point.  On failure (PK_ERR_CUDA: no peer access)
 *                  the scan simply runs without sharing.
 * Ranks must prepare the same queries and execute in lock-step for entries to be accepted; nothing else is required. */
int pk_fullscan_theta_export(pk_store* s, void* handle64);
void* pk_fullscan_theta_ptr(pk_store* s);
int pk_fullscan_theta_peers(pk_store* s, const void* handles, int n_ranks, int rank, const void* const* d_ptrs);
int pk_fullscan_execute_shortlists(pk_store* s);
int pk_fullscan_shortlist_keys(pk_store* s, uint64_t* d_short);
int pk_fullscan_execute_finish(pk_store* s, const uint64_t* d_all_short, int n_lists, int rank, int n_ranks);

int pk_sync(pk_store* s);
/* The CUDA stream (cudaStream_t) every kernel of this store is launched on. */
void* pk_stream(pk_store* s);

/* ---- introspection (bench / tests) ----------------------------------------------------------- */

enum {
    PK_STAT_KERNEL_LAUNCHES = 0,   /* kernels launched by this store since creation */
    PK_STAT_PAIRS_SCORED = 1,      /* (query, candidate) pairs of the last scoring call */
    PK_STAT_STORE_BYTES = 2,       /* bytes of the packed device record store (whole world) */
    PK_STAT_SCAN_KERNEL_NS = 3,    /* device time of the scan kernel(s) of the last scoring call (CUDA events) */
    PK_STAT_SCAN_LAUNCHES = 4,     /* scan kernel launches of the last scoring call */
    PK_STAT_GROUPS = 5,            /* query groups of the last scoring call */
    PK_STAT_SHARD_RECORD_BYTES = 6,/* bytes of the packed records inside the candidate shard */
    PK_STAT_H2D_BYTES = 7,         /* host-to-device bytes copied by this store since creation */
    PK_STAT_D2H_BYTES = 8,         /* device-to-host result bytes copied since creation */
    PK_STAT_PREFILTER = 9,         /* 1 when the prepared full scan runs the fp32 pre-filter + exact re-score pipeline */
    PK_STAT_FALLBACK_QUERIES = 10, /* queries of the last fetched full scan whose shortlist was not provably complete
                                      (they were re-run by the exact scan on every pair) */
    PK_STAT_RESCORE_NS = 11,       /* device time of the exact re-score + fallback kernels of the last full scan */
    PK_STAT_SHARD_LEAN_BYTES = 12, /* bytes of the lean fp32 records inside the candidate shard */
    PK_STAT_SHORTLIST = 13,        /* shortlist length Ks of the prepared full scan (0 = no pre-filter) */
    PK_STAT_MARGIN_E9 = 14,        /* the pre-filter's error bound on a FAS value, times 1e9 */
    PK_STAT_EDGE_CACHE = 15,       /* adjacency entries held by a valid per-edge FAS cache (0 = none) */
    PK_STAT_PREP_NS = 16,          /* device time of the query-table build of the last two-stage full scan */
    PK_STAT_MERGE_NS = 17          /* device time of its merge / proof kernels (instances, gathered shortlists, final) */
};
int64_t pk_stat(pk_store* s, int which);

/* Test hook of the two-stage full scan: the pre-filter's fp32 scores and the exact kernel's scores of
 * nq <= 32 queries against the store rows [row_begin,row_end), both as [row][nq].  Their difference must
 * stay below the bound PK_STAT_MARGIN_E9 reports; the parity tests assert it. */
int pk_debug_prefilter(pk_store* s, int nq, const int32_t* q_uid, int64_t row_begin, int64_t row_end, float* out_approx,
                       float* out_exact);

/* ---- binary profile store on disk ------------------------------------------------------
 * Replaces, for callers of this ABI, the CSV parse of the reference's load_users_encoded
 * (src/user_loader.cpp:10-96: single-threaded getline / stoi per cell, capped at 100 000 users, :34) and the
 * adjacency.csv reload (src/graph_builder.cpp:39-75) with one mmap: the file holds the arrays of a pk_world
 * (header + section table + 19 sections, 64-byte aligned, little-endian, checksummed header).
 *   pk_world_save   writes w to path (through path.tmp + rename)
 *   pk_world_open   maps path read-only; *view's pointers point into the mapping and stay valid until pk_world_close
 *                   (pass view to pk_store_load_world, which copies what it keeps).  A file that is not a world
 *                   file, is truncated, or whose section sizes contradict its counts fails with PK_ERR_ARG.
 *   pk_world_close  unmaps; NULL is a no-op
 * No CUDA device is needed for these three calls. */
typedef struct pk_world_file pk_world_file;
int pk_world_save(const pk_world* w, const char* path);
int pk_world_open(const char* path, pk_world_file** out, pk_world* view);
void pk_world_close(pk_world_file* f);

#ifdef __cplusplus
}
#endif
#endif
