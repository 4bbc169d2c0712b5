// pk_layout.h -- data layout shared by the host planner and the sm_100a kernels.
//
// PACKED PROFILE RECORD (one per user, 16-byte aligned, addressed by rec_off[row] in uint4 units).
// It is the device form of the reference's UserProfile (include/user_profile.h:10-20) with the
// per-(user,column) constants of tfidf_cosine_for_column (src/recommender.cpp:68-117) -- the
// weights tf*idf and the L2 norm -- precomputed once instead of on every call.
//
//   word  0      uid
//         1      public_flag          (<0 = missing,  src/recommender_similarity.cpp:38)
//         2      gender               (<0 = missing,  :45)
//         3      completion_percentage(<=0 = missing, :52)
//         4      age                  (<=0 = missing, :61)
//         5..7   region_parts[3]      (<0 = missing,  :70-72)
//         8      n_clubs     9 n_friends     10 n_cells (non-empty text columns)    11 n_tok
//         12,13  64-bit mask of non-empty text columns
//         14     number of region parts >= 0          15 reserved
//         16,17  sqrt((double)n_clubs)      18,19  sqrt((double)n_friends)
//   then  clubs[n_clubs]      uint32, padded to a multiple of 4 words
//         friends[n_friends]  uint32, padded to a multiple of 4 words
//         cells[n_cells]      4 words each: {column t, end token index (exclusive), norm (double)}
//                             ascending t; norm = sqrt(sum w^2) summed in ascending token order
//         tok_key[n_tok]      (t << 26) | token id, ascending inside a cell, padded to 4 words
//         tok_w[n_tok]        double w = (double)tf * (double)idf   (idf 1.0f when the token is
//                             missing from the column's table; w = tf when the column has no table)
//
// LEAN RECORD (one per user, 16-byte aligned, addressed by lrec_off[row] in uint4 units): the fp32 stream
// the pre-filter of the full scan reads (fas_prefilter_kernel).  The tf-idf weights are stored already
// divided by the cell's L2 norm, so a dot product of two lean cells IS the cosine (approximately).
//
//   word  0 uid   1 public   2 gender   3 completion   4 age   5..7 region_parts
//         8 n_clubs   9 n_friends   10 n_tok   11 number of region parts >= 0
//         12,13 64-bit mask of non-empty text columns   14 rsqrt(n_clubs) (float)   15 rsqrt(n_friends) (float)
//   then  clubs[n_clubs], friends[n_friends], tok_key[n_tok], tok_wn[n_tok] (float), each padded to 4 words
#pragma once
#include <stdint.h>

namespace pk {

constexpr int kHdrWords = 20;
constexpr int kHdrU4 = 5;
constexpr int kMaxT = 64;
constexpr int kTokBits = 26;            // token ids < 2^26
constexpr int kLanes = 32;              // queries per group = lanes of a warp
constexpr int kMaxK = 256;              // top-k list capacity of the full scan's fused shared-memory lists
constexpr int kMaxListK = 8192;         // top-k limit of the list modes (sort-based selection, pk_scan.cuh merge_topk_kernel)
constexpr int kNumFixed = 7;
constexpr int kMaxFields = kNumFixed + kMaxT;

__host__ __device__ inline uint32_t pad4(uint32_t x) { return (x + 3u) & ~3u; }

// size of a packed record in uint4 units
__host__ __device__ inline uint32_t record_u4(uint32_t n_clubs, uint32_t n_friends, uint32_t n_cells, uint32_t n_tok) {
    return kHdrU4 + pad4(n_clubs) / 4 + pad4(n_friends) / 4 + n_cells + pad4(n_tok) / 4 + (n_tok + 1) / 2;
}

constexpr int kLeanHdrWords = 16;
__host__ __device__ inline uint32_t lean_record_u4(uint32_t n_clubs, uint32_t n_friends, uint32_t n_tok) {
    return kLeanHdrWords / 4 + pad4(n_clubs) / 4 + pad4(n_friends) / 4 + 2 * (pad4(n_tok) / 4);
}

// sort key of the reference comparator (score desc, id asc; src/recommender_graph.cpp:97-100):
// larger key = better.  Scores are non-negative floats so their bits order like the values.
__host__ __device__ inline uint64_t make_key(uint32_t score_bits, int32_t id) {
    uint32_t low = ~((uint32_t)id ^ 0x80000000u);
    return ((uint64_t)score_bits << 33) | ((uint64_t)low << 1) | 1ull;
}
__host__ __device__ inline uint32_t key_score_bits(uint64_t k) { return (uint32_t)(k >> 33); }
__host__ __device__ inline int32_t key_id(uint64_t k) { return (int32_t)((~(uint32_t)(k >> 1)) ^ 0x80000000u); }

// Normalizer-derived constants (host-computed with the host libm so that every value that depends
// only on a discrete similarity is bit-identical to the reference's sigmoid(z(s))).
struct NormTables {
    double sig0[kMaxFields];   // sigmoid(z_f(0.0))
    double sig1[kMaxFields];   // sigmoid(z_f(1.0))
    double mean[kMaxFields];
    double sd[kMaxFields];
    double ftab[kMaxFields + 1];   // F = used / total_possible
    double region[64];             // [(a_cnt*4 + b_cnt)*4 + matches] -> sigmoid(z_region((float)s))
    uint8_t use_norm[kMaxFields];  // has entry && sd > 0  (src/recommender_similarity.cpp:30,107)
    uint8_t pad_[1];
    int32_t total_possible;        // 7 + T
    int32_t dim_comp, dim_age;     // completion / age tables are [dim][dim], index [cand][query]
};

// fp32 constants of the pre-filter.  sigma(z_f(s)) = 1 / (1 + 2^(a[f]*s + b[f])).
struct LeanNorm {
    float4 fld[kMaxFields];        // {a, b, sigmoid(z(0)), sigmoid(z(1))} per field
    float region[64];              // as NormTables::region
    float lut[kMaxT / 8][256];     // lut[j][bits] = sum of sigmoid(z(0)) over the text columns 8j+i with bit i set
    float inv_total;               // 1 / total_possible
    float margin;                  // bound on |pre-filter score - exact score|
    int32_t dim_comp, dim_age;
    // header-only upper bound of the pre-filter (pk_prefilter.cuh): the largest sigma a jointly filled text column /
    // a clubs / a friends field can contribute, rounded up
    float ub_text, ub_clubs, ub_friends, pad_;
};

// One group of <= 32 query users = the lanes of the scoring warps.
struct GroupDesc {
    uint32_t arena_off;     // byte offset of the group's tables in the arena (16-byte aligned)
    uint32_t arena_bytes;   // bytes to stage into shared memory
    int32_t n_lanes;
    int32_t in_smem;        // tables fit the shared-memory budget
    int32_t log_sc, log_sf, log_sa, log_st;        // log2 sizes of clubs / friends / adjacency / text tables
    uint32_t off_nq, off_clubs, off_friends, off_adj, off_text, off_vals;   // byte offsets inside the arena block
    uint32_t n_vals;
    // the block is laid out [lean-only | exclusion table | exact-only]: the pre-filter stages
    // [0, lean_end), the exact kernel [exact_begin, arena_bytes)
    // lean-only: the pre-filter's token table -- a 2-choice cuckoo table of 2^log_cb buckets x 2 slots {key, entry + 1}
    // (off_text2; probed with two independent 128-bit loads, no probe loop), the entries {query mask, first value index}
    // (off_base) and the normalised fp32 weights (off_vals32)
    uint32_t off_text2, off_base, off_vals32;
    int32_t log_cb;
    // lean-only: clubs and friends of the group's queries in ONE cuckoo table {id | set << 31, query mask}
    // (set 0 = clubs, 1 = friends; ids are < 2^31), 2^log_cs buckets x 2 slots
    uint32_t off_sets2;
    int32_t log_cs;
    uint32_t lean_end, exact_begin;
    int32_t lean_in_smem;
    int32_t adj_begin[kLanes + 1];   // slice of the per-batch exclusion id array for each lane
    int32_t row[kLanes];             // store row of each lane's user, -1 = unused lane
    int32_t qidx[kLanes];            // caller-side query index of each lane
};

struct WorkItem {
    int32_t group;
    int32_t cand_begin, cand_end;   // slice of the candidate list (or of the shard's row range)
    int32_t out_slot;               // top-k mode: index of the [32][Kp] key block; store mode: first output row
};

}  // namespace pk
