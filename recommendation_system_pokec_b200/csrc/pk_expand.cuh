// pk_expand.cuh -- friend-of-friend candidate expansion on the device (batched list modes).
//
// The reference enumerates candidates by walking the adjacency in order, keeping first occurrences and stopping at
// candidate_limit:
//   gather_candidates_local (src/recommender_graph.cpp:10-31)  friend, then that friend's friends, for every friend
//   recommend_collaborative (src/recommender_graph.cpp:114-125) the friends' friends only
// Both are "the first `limit` first-occurrences of a sequence", so a warp can take the sequence 32 elements at a
// time: probe a hash set of what earlier chunks inserted, elect one lane per distinct new id inside the chunk
// (__match_any_sync), give the elected lanes consecutive output positions in lane order (ballot + popc) and insert
// their ids.  The set lives in a per-warp scratch table in global memory (L2) tagged with an epoch, so it is never
// cleared between queries.  The same primitive numbers the distinct friends that have a profile (the columns of the
// collaborative FAS matrix).  One warp per query user; the adjacency CSR and the uid -> row maps are device-resident.
#pragma once
#include "pk_layout.h"

namespace pk {

struct DevGraph {
    const int64_t* adj_ptr;          // [m + 1]
    const int32_t* adj;              // neighbour uids, original order
    const int32_t* uid;              // [n] row -> uid
    // uid -> adjacency row / profile row: direct arrays over [uid_lo, uid_lo + range) when the uid range is dense
    // enough, else binary search over the ascending (uid, row) pairs
    const int32_t* adj_direct;
    const int32_t* row_direct;
    int32_t uid_lo;
    int64_t range;
    const int32_t* adj_sorted_uid;   // [m]
    const int32_t* adj_sorted_row;
    const int32_t* uid_sorted;       // [n]
    const int32_t* uid_sorted_row;
    int m, n;
};

__device__ __forceinline__ int dg_search(const int32_t* keys, const int32_t* vals, int cnt, int32_t u) {
    int lo = 0, hi = cnt;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (__ldg(keys + mid) < u) lo = mid + 1; else hi = mid; }
    return (lo < cnt && __ldg(keys + lo) == u) ? __ldg(vals + lo) : -1;
}
__device__ __forceinline__ int dg_adj_row(const DevGraph& G, int32_t u) {
    if (G.adj_direct) { const int64_t i = (int64_t)u - G.uid_lo; return (i >= 0 && i < G.range) ? __ldg(G.adj_direct + i) : -1; }
    return dg_search(G.adj_sorted_uid, G.adj_sorted_row, G.m, u);
}
__device__ __forceinline__ int dg_row(const DevGraph& G, int32_t u) {
    if (G.row_direct) { const int64_t i = (int64_t)u - G.uid_lo; return (i >= 0 && i < G.range) ? __ldg(G.row_direct + i) : -1; }
    return dg_search(G.uid_sorted, G.uid_sorted_row, G.n, u);
}

// ---- per-warp scratch set: slot = [epoch 8 | value 24 | key 32]; a slot belongs to the current query iff its epoch matches
constexpr int kExpWarps = 8;                     // warps per CTA of the expansion kernel
constexpr unsigned kExpFull = 0xffffffffu;

__device__ __forceinline__ uint32_t exp_hash(uint32_t key, uint32_t mask) { return (key * 0x9E3779B1u) >> 7 & mask; }

// slot index of `key`, or -1
__device__ __forceinline__ int set_find(const unsigned long long* tab, uint32_t mask, uint32_t epoch, uint32_t key) {
    uint32_t s = exp_hash(key, mask);
    for (;;) {
        const unsigned long long v = __ldcg(tab + s);
        if ((uint32_t)(v >> 56) != epoch) return -1;
        if ((uint32_t)v == key) return (int)s;
        s = (s + 1) & mask;
    }
}
// inserts a key that is known to be absent (one inserting lane per distinct key); returns its slot
__device__ __forceinline__ int set_insert(unsigned long long* tab, uint32_t mask, uint32_t epoch, uint32_t key, uint32_t value) {
    const unsigned long long nw = ((unsigned long long)epoch << 56) | ((unsigned long long)(value & 0xffffffu) << 32) | key;
    uint32_t s = exp_hash(key, mask);
    for (;;) {
        unsigned long long v = __ldcg(tab + s);
        while ((uint32_t)(v >> 56) != epoch) {
            const unsigned long long old = atomicCAS(tab + s, v, nw);
            if (old == v) return (int)s;
            v = old;
        }
        s = (s + 1) & mask;
    }
}

struct ExpandParams {
    DevGraph G;
    const int32_t* q_uid;
    int nq;
    int mode;                        // PK_MODE_GRAPH: gather_candidates_local; PK_MODE_COLLAB / PK_MODE_CLUBS: friends' friends
    int limit;                       // >= 1
    int raw;                         // 1 = emit the raw expansion (uids) without the profile / friend filters (pk_candidates_batch)
    int slots_only;                  // 1 = only number the friends (recommend_clubs_collab enumerates edges, not candidates)
    unsigned long long* scratch;     // [warps][tab_cap] persistent across launches
    uint32_t* epoch;                 // [warps] last epoch used
    uint32_t tab_cap;                // power of two >= 2 (limit + 32) and >= 2 (max friends of a query) -- see the host
    int* next_query;                 // work queue (zeroed per launch)
    int cand_cap;                    // per-query capacity of cand_out (= limit)
    // outputs
    int32_t* q_row;                  // [nq] profile row of the user, -1 = none (no result)
    int32_t* cand_out;               // [nq][cand_cap] kept candidates in the reference's order: rows (raw: uids)
    int32_t* cand_cnt;               // [nq]
    // collaborative / clubs: the distinct friends that have a profile, numbered in adjacency order
    int32_t* fr_slot;                // [E] per adjacency entry of a queried user: slot, -1 = no profile
    int32_t* slot_row;               // [E] per slot (stored at the user's adjacency offset): profile row
    int32_t* n_slots;                // [nq]
    int32_t* q_adj_off;              // [nq] offset of the user's adjacency row in `adj` (-1 = no row)
    int32_t* q_deg;                  // [nq]
};

__global__ void __launch_bounds__(kExpWarps * 32) expand_kernel(const ExpandParams P) {
    const int lane = threadIdx.x & 31;
    const int wslot = blockIdx.x * kExpWarps + (threadIdx.x >> 5);
    unsigned long long* tab = P.scratch + (size_t)wslot * P.tab_cap;
    const uint32_t mask = P.tab_cap - 1u;
    const uint32_t lane_lt = (1u << lane) - 1u;
    const DevGraph& G = P.G;
    uint32_t epoch = P.epoch[wslot];
    auto next_epoch = [&]() {
        if (++epoch > 255u) {                       // wrapped: every old tag would alias -- clear the table once
            for (uint32_t i = lane; i < P.tab_cap; i += 32) tab[i] = 0ull;
            __syncwarp();
            epoch = 1u;
        }
    };
    for (;;) {
        int q = 0;
        if (lane == 0) q = atomicAdd(P.next_query, 1);
        q = __shfl_sync(kExpFull, q, 0);
        if (q >= P.nq) break;
        const int32_t user = P.q_uid[q];
        const int urow = dg_row(G, user);
        const int arow = dg_adj_row(G, user);
        const int64_t ub = arow >= 0 ? G.adj_ptr[arow] : 0;
        const int deg = arow >= 0 ? (int)(G.adj_ptr[arow + 1] - ub) : 0;
        if (lane == 0) {
            P.q_row[q] = urow;
            P.cand_cnt[q] = 0;
            if (P.n_slots) { P.n_slots[q] = 0; P.q_adj_off[q] = arow >= 0 ? (int32_t)ub : -1; P.q_deg[q] = deg; }
        }
        // src/recommender_graph.cpp:40,130, src/recommender_clubs.cpp:16: no profile, no result (the raw lists of
        // pk_candidates do not depend on the profile)
        if ((urow < 0 && !P.raw) || deg == 0) continue;

        // ---- the distinct friends that have a profile, in adjacency order (collaborative / clubs) ----
        if (P.n_slots && !P.raw) {
            next_epoch();
            int ns = 0;
            for (int i0 = 0; i0 < deg; i0 += 32) {
                const int i = i0 + lane;
                int32_t f = 0;
                int rf = -1;
                if (i < deg) { f = __ldg(G.adj + ub + i); rf = dg_row(G, f); }
                const bool valid = rf >= 0;                                   // src/recommender_graph.cpp:134
                int sl = valid ? set_find(tab, mask, epoch, (uint32_t)f) : -1;
                int slot = sl >= 0 ? (int)((__ldcg(tab + sl) >> 32) & 0xffffffu) : -1;
                const bool fresh = valid && sl < 0;
                const uint32_t fresh_m = __ballot_sync(kExpFull, fresh);
                // one lane per distinct new friend: the lowest lane holding that uid
                const uint32_t same = __match_any_sync(kExpFull, fresh ? f : (int32_t)(0x80000000u | (uint32_t)lane)) & fresh_m;
                const bool leader = fresh && (__ffs(same) - 1) == lane;
                const uint32_t lead_m = __ballot_sync(kExpFull, leader);
                const int pos = ns + __popc(lead_m & lane_lt);
                if (leader) {
                    set_insert(tab, mask, epoch, (uint32_t)f, (uint32_t)pos);
                    P.slot_row[ub + pos] = rf;
                }
                const int lpos = __shfl_sync(kExpFull, pos, fresh ? (__ffs(same) - 1) : 0);
                if (fresh) slot = lpos;
                if (i < deg) P.fr_slot[ub + i] = slot;
                ns += __popc(lead_m);
                __syncwarp();
            }
            if (lane == 0) P.n_slots[q] = ns;
        }

        if (P.slots_only) continue;
        // ---- candidate expansion: first `limit` first-occurrences of the reference's visiting sequence ----
        next_epoch();
        int count = 0;
        int32_t* out = P.cand_out + (size_t)q * P.cand_cap;
        const bool graph = P.mode == PK_MODE_GRAPH;
        for (int fi = 0; fi < deg && count < P.limit; ++fi) {
            const int32_t f = __ldg(G.adj + ub + fi);
            if (graph && f == user) continue;                                 // src/recommender_graph.cpp:18
            const int frow = dg_adj_row(G, f);
            const int64_t fb = frow >= 0 ? G.adj_ptr[frow] : 0;
            const int fdeg = frow >= 0 ? (int)(G.adj_ptr[frow + 1] - fb) : 0;
            // graph: the friend itself, then its row (:19-29); collaborative: the row only (:116-124)
            const int total = fdeg + (graph ? 1 : 0);
            for (int e0 = 0; e0 < total && count < P.limit; e0 += 32) {
                const int e = e0 + lane;
                int32_t x = 0;
                bool valid = e < total;
                if (valid) {
                    x = (graph && e == 0) ? f : __ldg(G.adj + fb + e - (graph ? 1 : 0));
                    valid = x != user;                                        // :23 / :120
                }
                const bool fresh = valid && set_find(tab, mask, epoch, (uint32_t)x) < 0;
                const uint32_t fresh_m = __ballot_sync(kExpFull, fresh);
                const uint32_t same = __match_any_sync(kExpFull, fresh ? x : (int32_t)(0x80000000u | (uint32_t)lane)) & fresh_m;
                const bool leader = fresh && (__ffs(same) - 1) == lane;
                const uint32_t lead_m = __ballot_sync(kExpFull, leader);
                const int pos = count + __popc(lead_m & lane_lt);
                if (leader && pos < P.limit) {
                    set_insert(tab, mask, epoch, (uint32_t)x, 0u);
                    out[pos] = x;
                }
                count += __popc(lead_m);
                __syncwarp();
            }
        }
        count = min(count, P.limit);
        if (P.raw) {
            if (lane == 0) P.cand_cnt[q] = count;
            continue;
        }
        // ---- filters: a profile must exist (:54 / :172); graph also drops the user's existing friends (:46-52) ----
        if (graph) {
            for (int i = lane; i < deg; i += 32) {
                const int sl = set_find(tab, mask, epoch, (uint32_t)__ldg(G.adj + ub + i));
                if (sl >= 0) atomicOr(tab + sl, 1ull << 32);                  // value bit 0 = "is a friend"
            }
            __syncwarp();
        }
        int kept = 0;
        for (int i0 = 0; i0 < count; i0 += 32) {
            const int i = i0 + lane;
            int rc = -1;
            if (i < count) {
                const int32_t c = out[i];
                rc = dg_row(G, c);
                if (graph && rc >= 0) {
                    const int sl = set_find(tab, mask, epoch, (uint32_t)c);
                    if (sl >= 0 && ((__ldcg(tab + sl) >> 32) & 1ull)) rc = -1;
                }
            }
            const uint32_t keep_m = __ballot_sync(kExpFull, rc >= 0);
            __syncwarp();                                                     // every lane has read out[i0 .. i0 + 32)
            if (rc >= 0) out[kept + __popc(keep_m & lane_lt)] = rc;           // in place: kept <= i0
            kept += __popc(keep_m);
            __syncwarp();
        }
        if (lane == 0) P.cand_cnt[q] = kept;
    }
    if (lane == 0) P.epoch[wslot] = epoch;
}

// ---- offsets: exclusive scan of per-query counts (single CTA; nq is a batch, not a world) ----
// in: cnt_a[nq] (candidates), cnt_b[nq] (slots, may be null).  out: cand_off[nq + 1]; pair_off[nq + 1] with per-query
// pair count = graph: cnt_a; collaborative: cnt_b (the w_f pairs) + cnt_a * cnt_b (the FAS(f, c) matrix)
__global__ void __launch_bounds__(1024) offsets_kernel(int nq, const int32_t* __restrict__ cnt_a, const int32_t* __restrict__ cnt_b,
                                                       int32_t* __restrict__ cand_off, int64_t* __restrict__ pair_off, int matrix_pairs) {
    __shared__ long long wsum_a[32], wsum_p[32];
    __shared__ long long run_a, run_p;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { run_a = 0; run_p = 0; }
    __syncthreads();
    for (int base = 0; base < nq; base += 1024) {
        const int q = base + tid;
        long long a = 0, p = 0;
        if (q < nq) {
            a = cnt_a[q];
            p = cnt_b ? (long long)cnt_b[q] * (matrix_pairs ? 1 + a : 1) : a;
        }
        long long ia = a, ip = p;
        for (int o = 1; o < 32; o <<= 1) {
            const long long va = __shfl_up_sync(0xffffffffu, ia, o), vp = __shfl_up_sync(0xffffffffu, ip, o);
            if (lane >= o) { ia += va; ip += vp; }
        }
        if (lane == 31) { wsum_a[warp] = ia; wsum_p[warp] = ip; }
        __syncthreads();
        long long oa = run_a, op = run_p;
        for (int w = 0; w < warp; ++w) { oa += wsum_a[w]; op += wsum_p[w]; }
        if (q < nq) { cand_off[q] = (int32_t)(oa + ia - a); pair_off[q] = op + ip - p; }
        __syncthreads();
        if (tid == 1023) { run_a = oa + ia; run_p = op + ip; }
        __syncthreads();
    }
    if (tid == 0) { cand_off[nq] = (int32_t)run_a; pair_off[nq] = run_p; }
}

// ---- pair emission: one CTA per query ----
struct EmitParams {
    int nq, mode, cand_cap;
    const int32_t* q_row;
    const int32_t* cand_out;     // [nq][cand_cap] candidate rows
    const int32_t* cand_cnt;
    const int32_t* cand_off;     // [nq + 1]
    const int64_t* pair_off;     // [nq + 1]
    const int32_t* n_slots;      // collaborative
    const int32_t* q_adj_off;
    const int32_t* slot_row;
    const int32_t* uid;          // row -> uid
    int32_t* a_row;              // [pairs]
    int32_t* b_row;
    int32_t* key_id;             // graph: [pairs] candidate uid of each pair; collaborative: [cands] candidate uid
    int32_t* cand_row_flat;      // collaborative: [cands] candidate rows, compacted (grouped by cost class)
    const uint32_t* cost;        // [n] per profile row: tokens + friends + clubs (what a pair's merges walk)
    int friend_major;            // matrix pair index = slot * nc + c (a warp holds one friend) or c * ns + slot
    int matrix_pairs;            // 0: the FAS(f, c) matrix goes to the group kernel (collab_groups_*), only the w_f pairs are emitted
};

__global__ void __launch_bounds__(256) emit_pairs_kernel(const EmitParams P) {
    const int q = blockIdx.x;
    const int nc = P.cand_cnt[q];
    const int32_t* cand = P.cand_out + (size_t)q * P.cand_cap;
    const int64_t p0 = P.pair_off[q];
    const int c0 = P.cand_off[q];
    const int urow = P.q_row[q];
    if (P.mode == PK_MODE_GRAPH) {
        for (int i = threadIdx.x; i < nc; i += blockDim.x) {
            const int rc = cand[i];
            P.a_row[p0 + i] = urow;
            P.b_row[p0 + i] = rc;
            P.key_id[p0 + i] = P.uid[rc];
        }
        return;
    }
    const int ns = P.n_slots[q];
    const int32_t* srow = P.slot_row + (P.q_adj_off[q] >= 0 ? P.q_adj_off[q] : 0);
    // The candidates of a user are scored in any order (only their keys matter afterwards), so they are grouped by
    // cost class (power-of-two bands of tokens + friends + clubs): the 32 lanes of a warp of the pair kernel then
    // walk records of similar length instead of all waiting for the longest one.
    __shared__ int cls_cnt[16], cls_base[16];
    if (threadIdx.x < 16) cls_cnt[threadIdx.x] = 0;
    __syncthreads();
    auto cls_of = [&](int rc) { const uint32_t c = __ldg(P.cost + rc) >> 4; return c ? min(15, 32 - __clz(c)) : 0; };
    for (int i = threadIdx.x; i < nc; i += blockDim.x) atomicAdd(&cls_cnt[cls_of(cand[i])], 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        int run = 0;
        for (int k = 0; k < 16; ++k) { cls_base[k] = run; run += cls_cnt[k]; cls_cnt[k] = 0; }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nc; i += blockDim.x) {
        const int rc = cand[i];
        const int k = cls_of(rc);
        const int pos = cls_base[k] + atomicAdd(&cls_cnt[k], 1);
        P.cand_row_flat[c0 + pos] = rc;
        P.key_id[c0 + pos] = P.uid[rc];
    }
    __syncthreads();
    // w_f = FAS(user, f) for every distinct friend with a profile, then FAS(f, c) friend-major: pair index =
    // slot * nc + c, so that a warp holds ONE friend (broadcast loads of the A side) and 32 candidates of one class
    for (int s = threadIdx.x; s < ns; s += blockDim.x) { P.a_row[p0 + s] = urow; P.b_row[p0 + s] = srow[s]; }
    const int64_t m0 = p0 + ns;
    const int64_t total = P.matrix_pairs ? (int64_t)nc * ns : 0;
    for (int64_t t = threadIdx.x; t < total; t += blockDim.x) {
        int sl, c;
        if (P.friend_major) { sl = (int)(t / nc); c = (int)(t - (int64_t)sl * nc); }
        else { c = (int)(t / ns); sl = (int)(t - (int64_t)c * ns); }
        P.a_row[m0 + t] = srow[sl];
        P.b_row[m0 + t] = P.cand_row_flat[c0 + c];
    }
}

// collaborative reduce on the pair layout above (src/recommender_graph.cpp:167-180,213): score(c) = sum over the
// user's friends IN ADJACENCY ORDER (duplicates counted again, friends without a profile skipped) of
// (double)w_f * (double)FAS(f, c), rounded to float once
struct CollabReduce2Params {
    int nq;
    const float* fas;            // pair scores
    const int64_t* pair_off;
    const int32_t* cand_off;
    const int32_t* n_slots;
    const int32_t* q_adj_off;
    const int32_t* q_deg;
    const int32_t* fr_slot;      // [E]
    const int32_t* key_id;       // [cands]
    uint64_t* out_keys;          // [cands]
    int friend_major;
};

__global__ void __launch_bounds__(128) collab_reduce2_kernel(const CollabReduce2Params P) {
    const int q = blockIdx.y;
    const int c0 = P.cand_off[q], nc = P.cand_off[q + 1] - c0;
    if (nc == 0) return;
    const int ns = P.n_slots[q], deg = P.q_deg[q];
    const float* w = P.fas + P.pair_off[q];
    const float* mat = w + ns;
    const int32_t* fs = P.fr_slot + (P.q_adj_off[q] >= 0 ? P.q_adj_off[q] : 0);
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < nc; c += gridDim.x * blockDim.x) {
        double score = 0.0;
        for (int i = 0; i < deg; ++i) {
            const int sl = fs[i];
            if (sl < 0) continue;
            score = __dadd_rn(score, __dmul_rn((double)w[sl], (double)mat[P.friend_major ? (int64_t)sl * nc + c : (int64_t)c * ns + sl]));
        }
        P.out_keys[c0 + c] = make_key(__float_as_uint((float)score), P.key_id[c0 + c]);
    }
}

// ---- clubs (src/recommender_clubs.cpp:23-64): every user's ordered edge stream, built on the device ----------
// 1-hop entries (every friend entry whose friend has a profile, :34-43), then 2-hop entries (for every such friend
// with an adjacency row, every neighbour != user that has a profile, :45-64), each with the indices of the FAS values
// it needs and the offset of its club contributions.  COUNT pass: sizes per user; FILL pass: the stream itself and
// the (A, B) pairs the lane-per-pair kernel scores (none when the per-edge FAS cache serves them).
struct ClubsPlanParams {
    DevGraph G;
    const int32_t* q_uid;
    int nq;
    const int32_t *q_row, *n_slots, *q_adj_off, *q_deg, *fr_slot, *slot_row;     // from expand_kernel
    const uint32_t* ncl;             // [n] clubs per profile row
    int cached;                      // FAS values come from the per-edge cache: indices are adjacency offsets
    int32_t* n_edges;                // COUNT out [nq]
    int32_t* n_two;                  // COUNT out [nq] 2-hop entries
    unsigned long long* n_con;       // COUNT out [nq] club contributions
    const int32_t* edge_off;         // FILL in [nq + 1]
    const int64_t* pair_off;         // FILL in [nq + 1]
    ClubEdge* edges;                 // FILL out
    uint32_t* con_off;               // FILL out, per edge, relative to the user
    int32_t *a_row, *b_row;          // FILL out (not cached)
};

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane) {
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(kExpFull, v, o); if (lane >= o) v += t; }
    return v;
}

template <bool FILL>
__global__ void __launch_bounds__(kExpWarps * 32) clubs_plan_kernel(const ClubsPlanParams P) {
    const int lane = threadIdx.x & 31;
    const uint32_t lane_lt = (1u << lane) - 1u;
    const DevGraph& G = P.G;
    for (int q = blockIdx.x * kExpWarps + (threadIdx.x >> 5); q < P.nq; q += gridDim.x * kExpWarps) {
        const int urow = P.q_row[q], deg = P.q_deg[q], ns = P.n_slots[q];
        const int64_t ub = P.q_adj_off[q];
        int ne = 0, n2 = 0;
        unsigned long long con = 0ull;
        if (urow >= 0 && ub >= 0 && deg > 0 && ns > 0) {
            const int32_t user = P.q_uid[q];
            const int64_t e0 = FILL ? P.edge_off[q] : 0, p0 = FILL ? P.pair_off[q] : 0;
            if (FILL && !P.cached)
                for (int s = lane; s < ns; s += 32) { P.a_row[p0 + s] = urow; P.b_row[p0 + s] = P.slot_row[ub + s]; }
            // 1-hop
            for (int i0 = 0; i0 < deg; i0 += 32) {
                const int i = i0 + lane;
                const int sl = (i < deg) ? P.fr_slot[ub + i] : -1;
                const bool valid = sl >= 0;
                const int tr = valid ? P.slot_row[ub + sl] : 0;
                const uint32_t c = valid ? __ldg(P.ncl + tr) : 0u;
                const uint32_t vm = __ballot_sync(kExpFull, valid);
                const uint32_t inc = warp_incl_scan(c, lane);
                if (FILL && valid) {
                    const int pos = ne + __popc(vm & lane_lt);
                    ClubEdge ed;
                    ed.w_idx = P.cached ? ub + i : p0 + sl; ed.s_idx = -1; ed.target_row = tr; ed.pad_ = 0;
                    P.edges[e0 + pos] = ed;
                    P.con_off[e0 + pos] = (uint32_t)min(con + inc - c, 0xffffffffull);
                }
                ne += __popc(vm);
                con += __shfl_sync(kExpFull, inc, 31);
            }
            // 2-hop
            for (int i = 0; i < deg; ++i) {
                const int sl = P.fr_slot[ub + i];
                if (sl < 0) continue;                                         // no profile (:49)
                const int32_t f = __ldg(G.adj + ub + i);
                const int fa = dg_adj_row(G, f);
                if (fa < 0) continue;                                         // no adjacency row (:47)
                const int64_t fb = G.adj_ptr[fa];
                const int fdeg = (int)(G.adj_ptr[fa + 1] - fb);
                const int frow = P.slot_row[ub + sl];
                for (int j0 = 0; j0 < fdeg; j0 += 32) {
                    const int j = j0 + lane;
                    int rx = -1;
                    if (j < fdeg) { const int32_t x = __ldg(G.adj + fb + j); if (x != user) rx = dg_row(G, x); }   // :53-55
                    const bool valid = rx >= 0;
                    const uint32_t c = valid ? __ldg(P.ncl + rx) : 0u;
                    const uint32_t vm = __ballot_sync(kExpFull, valid);
                    const uint32_t inc = warp_incl_scan(c, lane);
                    if (FILL && valid) {
                        const int k2 = n2 + __popc(vm & lane_lt);
                        ClubEdge ed;
                        ed.w_idx = P.cached ? ub + i : p0 + sl;
                        ed.s_idx = P.cached ? fb + j : p0 + ns + k2;
                        ed.target_row = rx; ed.pad_ = 0;
                        P.edges[e0 + ne + k2 - n2 + 0] = ed;
                        P.con_off[e0 + ne + k2 - n2] = (uint32_t)min(con + inc - c, 0xffffffffull);
                        if (!P.cached) { P.a_row[p0 + ns + k2] = frow; P.b_row[p0 + ns + k2] = rx; }
                    }
                    const int add = __popc(vm);
                    ne += add; n2 += add;
                    con += __shfl_sync(kExpFull, inc, 31);
                }
            }
        }
        if (!FILL && lane == 0) { P.n_edges[q] = ne; P.n_two[q] = n2; P.n_con[q] = con; }
    }
}

// offsets of the clubs plan: edges, pairs (n_slots + 2-hop entries per user unless cached) and the per-user club
// table slices (power of two >= 2 x contributions, as the scatter kernel expects)
__global__ void __launch_bounds__(1024) clubs_offsets_kernel(int nq, int cached, const int32_t* __restrict__ n_edges,
                                                             const int32_t* __restrict__ n_two, const int32_t* __restrict__ n_slots,
                                                             const unsigned long long* __restrict__ n_con, int32_t* __restrict__ edge_off,
                                                             int64_t* __restrict__ pair_off, int32_t* __restrict__ slot_begin,
                                                             int64_t* __restrict__ totals) {
    __shared__ long long ws[3][32];
    __shared__ long long run[3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 3) run[tid] = 0;
    __syncthreads();
    for (int base = 0; base < nq; base += 1024) {
        const int q = base + tid;
        long long v[3] = {0, 0, 0};
        if (q < nq) {
            v[0] = n_edges[q];
            v[1] = (cached || n_edges[q] == 0) ? 0 : (long long)n_slots[q] + n_two[q];
            unsigned long long c = n_con[q] < 1ull ? 1ull : n_con[q], cap = 2;
            while (cap < 2 * c) cap <<= 1;
            v[2] = (long long)cap;
        }
        long long inc[3] = {v[0], v[1], v[2]};
        for (int o = 1; o < 32; o <<= 1)
            for (int k = 0; k < 3; ++k) { const long long t = __shfl_up_sync(0xffffffffu, inc[k], o); if (lane >= o) inc[k] += t; }
        if (lane == 31) for (int k = 0; k < 3; ++k) ws[k][warp] = inc[k];
        __syncthreads();
        long long off[3] = {run[0], run[1], run[2]};
        for (int w = 0; w < warp; ++w) for (int k = 0; k < 3; ++k) off[k] += ws[k][w];
        if (q < nq) {
            edge_off[q] = (int32_t)min(off[0] + inc[0] - v[0], 0x7fffffffll);
            pair_off[q] = off[1] + inc[1] - v[1];
            slot_begin[q] = (int32_t)min(off[2] + inc[2] - v[2], 0x7fffffffll);
        }
        __syncthreads();
        if (tid == 1023) for (int k = 0; k < 3; ++k) run[k] = off[k] + inc[k];
        __syncthreads();
    }
    if (tid == 0) {
        edge_off[nq] = (int32_t)min(run[0], 0x7fffffffll); pair_off[nq] = run[1]; slot_begin[nq] = (int32_t)min(run[2], 0x7fffffffll);
        totals[0] = run[0]; totals[1] = run[1]; totals[2] = run[2];
    }
}

// ---- the FAS(f, c) matrix of recommend_collaborative through fas_scan_kernel: group planning on the device ----
// Per query, every block of 32 distinct friends is a group of the exact scan kernel (lane = friend) and the query's
// candidates are streamed past it once: 32 pairs share every candidate record and every table probe, which the
// lane-per-pair kernel cannot do.  What the host planner does for single-user calls (pk_host.cu: layout_group,
// collab_prepare) is done here for a whole batch: group descriptors, arena offsets, work items.
__host__ __device__ inline int pk_log2_table(unsigned long long entries) {       // smallest L >= 1 with 2^L >= 2 * entries
    int L = 1;
    while ((1ull << L) < 2ull * entries) ++L;
    return L;
}
__host__ __device__ inline uint32_t pk_align16(uint32_t x) { return (x + 15u) & ~15u; }

// table layout of a group of the EXACT kernel without exclusion ids (the layout pk_host.cu:layout_group produces with
// with_lean = false and no exclusion ids)
__host__ __device__ inline void pk_layout_exact_group(GroupDesc& g, unsigned long long ncl, unsigned long long nfr,
                                                      unsigned long long ntok, int T) {
    g.log_sc = pk_log2_table(ncl); g.log_sf = pk_log2_table(nfr); g.log_sa = pk_log2_table(0); g.log_st = pk_log2_table(ntok);
    uint32_t off = 0;
    g.off_text2 = g.off_base = g.off_vals32 = g.off_sets2 = 0; g.log_cb = g.log_cs = 0;
    g.exact_begin = off;
    g.off_adj = off; off += (8u << g.log_sa);
    off = pk_align16(off);
    g.lean_end = 0u;
    g.off_clubs = off; off += (8u << g.log_sc);
    g.off_friends = off; off += (8u << g.log_sf);
    off = pk_align16(off);
    g.off_nq = off; off += pk_align16((uint32_t)T * 32u * 8u);
    g.off_text = off; off += (16u << g.log_st);
    g.off_vals = off; off += pk_align16((uint32_t)ntok * 8u);
    g.n_vals = (uint32_t)ntok;
    g.arena_bytes = pk_align16(off);
}

struct CollabGroupsParams {
    int nq, T;
    uint32_t smem_budget;            // largest table block the scan kernel can stage in shared memory
    const int32_t *cand_cnt, *cand_off, *n_slots, *q_adj_off, *slot_row;
    const uint32_t *ncl, *nfr, *ntok;    // [n] per profile row
    int32_t* group_off;              // [nq + 1] first group of each query
    // per group
    GroupDesc* groups;
    WorkItem* items;
    unsigned long long* totals;      // {groups, arena bytes, output rows, max shared-memory block, groups outside shared memory}
};

// pass 1 (one CTA): groups per query -> group_off
__global__ void __launch_bounds__(1024) collab_groups_count_kernel(const CollabGroupsParams P) {
    __shared__ int wsum[32];
    __shared__ int run;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) run = 0;
    __syncthreads();
    for (int base = 0; base < P.nq; base += 1024) {
        const int q = base + tid;
        int v = 0;
        if (q < P.nq && P.cand_cnt[q] > 0) v = (P.n_slots[q] + 31) >> 5;
        int inc = v;
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        int off = run;
        for (int w = 0; w < warp; ++w) off += wsum[w];
        if (q < P.nq) P.group_off[q] = off + inc - v;
        __syncthreads();
        if (tid == 1023) run = off + inc;
        __syncthreads();
    }
    if (tid == 0) { P.group_off[P.nq] = run; P.totals[0] = (unsigned long long)run; }
}

// pass 2 (one CTA, sequential over the groups in chunks of 1024 threads): descriptors, arena offsets, work items.
// The running offsets are carried across the chunks by a block scan.
__global__ void __launch_bounds__(1024) collab_groups_fill_kernel(const CollabGroupsParams P) {
    __shared__ unsigned long long ws_a[32], ws_r[32];
    __shared__ unsigned long long run_a, run_r, max_sm, n_gl;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { run_a = 0; run_r = 0; max_sm = 0; n_gl = 0; }
    __syncthreads();
    const int G = P.group_off[P.nq];
    for (int base = 0; base < G; base += 1024) {
        const int g = base + tid;
        GroupDesc gd;
        unsigned long long a = 0, r = 0;
        int q = 0;
        if (g < G) {
            int lo = 0, hi = P.nq;                      // the query whose group range holds g
            while (lo < hi) { const int mid = (lo + hi) >> 1; if (P.group_off[mid + 1] <= g) lo = mid + 1; else hi = mid; }
            q = lo;
            const int b = g - P.group_off[q];
            const int ns = P.n_slots[q];
            const int32_t* srow = P.slot_row + P.q_adj_off[q];
            memset(&gd, 0, sizeof(gd));
            gd.n_lanes = min(32, ns - 32 * b);
            unsigned long long ncl = 0, nfr = 0, ntok = 0;
            for (int l = 0; l < 32; ++l) {
                int row = -1;
                if (l < gd.n_lanes) { row = srow[32 * b + l]; ncl += P.ncl[row]; nfr += P.nfr[row]; ntok += P.ntok[row]; }
                gd.row[l] = row; gd.qidx[l] = q;
            }
            pk_layout_exact_group(gd, ncl, nfr, ntok, P.T);
            gd.in_smem = gd.arena_bytes <= P.smem_budget ? 1 : 0;
            a = gd.arena_bytes;
            r = (unsigned long long)P.cand_cnt[q];
            if (gd.in_smem) atomicMax(&max_sm, (unsigned long long)gd.arena_bytes); else atomicAdd(&n_gl, 1ull);
        }
        unsigned long long ia = a, ir = r;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long ta = __shfl_up_sync(0xffffffffu, ia, o), tr = __shfl_up_sync(0xffffffffu, ir, o);
            if (lane >= o) { ia += ta; ir += tr; }
        }
        if (lane == 31) { ws_a[warp] = ia; ws_r[warp] = ir; }
        __syncthreads();
        unsigned long long oa = run_a, orr = run_r;
        for (int w = 0; w < warp; ++w) { oa += ws_a[w]; orr += ws_r[w]; }
        if (g < G) {
            gd.arena_off = (uint32_t)min(oa + ia - a, 0xffffffffull);
            P.groups[g] = gd;
            WorkItem it;
            it.group = g; it.cand_begin = P.cand_off[q]; it.cand_end = P.cand_off[q] + P.cand_cnt[q];
            it.out_slot = (int32_t)min(orr + ir - r, 0x7fffffffull);
            P.items[g] = it;
        }
        __syncthreads();
        if (tid == 1023) { run_a = oa + ia; run_r = orr + ir; }
        __syncthreads();
    }
    if (tid == 0) { P.totals[1] = run_a; P.totals[2] = run_r; P.totals[3] = max_sm; P.totals[4] = n_gl; }
}

// reduce on the group kernel's [row][32] output: FAS(f, c) of slot sl sits at (out_slot(group of sl) + c) * 32 + sl % 32
struct CollabReduce3Params {
    int nq;
    const float* w;              // pair scores: w_f of slot sl of query q at pair_off[q] + sl
    const float* mat;            // store-mode output of fas_scan_kernel
    const int64_t* pair_off;
    const int32_t *cand_off, *n_slots, *q_adj_off, *q_deg, *fr_slot, *group_off, *key_id;
    const WorkItem* items;
    uint64_t* out_keys;
};

__global__ void __launch_bounds__(128) collab_reduce3_kernel(const CollabReduce3Params P) {
    const int q = blockIdx.y;
    const int c0 = P.cand_off[q], nc = P.cand_off[q + 1] - c0;
    if (nc == 0) return;
    const int deg = P.q_deg[q], g0 = P.group_off[q];
    const float* w = P.w + P.pair_off[q];
    const int32_t* fs = P.fr_slot + (P.q_adj_off[q] >= 0 ? P.q_adj_off[q] : 0);
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < nc; c += gridDim.x * blockDim.x) {
        double score = 0.0;
        for (int i = 0; i < deg; ++i) {
            const int sl = fs[i];
            if (sl < 0) continue;
            const int64_t row = (int64_t)P.items[g0 + (sl >> 5)].out_slot + c;
            score = __dadd_rn(score, __dmul_rn((double)w[sl], (double)P.mat[row * 32 + (sl & 31)]));
        }
        P.out_keys[c0 + c] = make_key(__float_as_uint((float)score), P.key_id[c0 + c]);
    }
}

}  // namespace pk
