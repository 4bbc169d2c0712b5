// pk_scan.cuh -- the scoring kernels.
//
// fas_scan_kernel is the device form of Recommender::profile_similarity
// (reference src/recommender_similarity.cpp:10-124) applied to (query group x candidate) tiles,
// with the top-k of src/recommender_graph.cpp:97-101 fused in.
//
// Mapping.  A group is <= 32 query users; lane l of every warp of the CTA owns query l.  A warp
// takes one candidate record at a time: the record's clubs / friends / tokens are read with
// coalesced loads (lane i owns element i) and every element is looked up ONCE in the group's
// inverted hash tables (element -> 32-bit mask of the queries that contain it), which live in
// shared memory.  Hits are redistributed with warp shuffles so that lane l accumulates the overlap
// count / tf-idf dot product of ITS query, then each lane finishes its own FAS in fp64 exactly as
// the reference does (float rounding of every per-field similarity, z-score, logistic, mean,
// fill factor, harmonic mean, final float).  Control flow is warp-uniform except for predication.
#pragma once
#include "pk_build.cuh"
#include "pk_layout.h"

namespace pk {

#ifndef PK_SCAN_THREADS
#define PK_SCAN_THREADS 1024
#endif
constexpr int kScanThreads = PK_SCAN_THREADS;
constexpr int kScanWarps = kScanThreads / 32;
constexpr int kPrepThreads = 256;
constexpr unsigned kFull = 0xffffffffu;

enum { EPI_TOPK = 0, EPI_STORE = 1 };

// ------------------------------------------------------------------------------------------------
// group tables: built in the global arena by group_prep_kernel
// ------------------------------------------------------------------------------------------------

// set table slot: {key, mask}; text table slot: {key, mask, base, 0}.  A slot is empty iff mask == 0.
__device__ __forceinline__ void table_insert(unsigned long long* slot0, uint32_t stride_u64, int log_size, uint32_t key,
                                             uint32_t bit) {
    const uint32_t msk = (1u << log_size) - 1u;
    uint32_t slot = hash_u32(key, log_size);
    const unsigned long long nw = (unsigned long long)key | ((unsigned long long)bit << 32);
    for (;;) {
        unsigned long long* p = slot0 + (size_t)slot * stride_u64;
        unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(p);
        if (cur == 0ull) {
            unsigned long long old = atomicCAS(p, 0ull, nw);
            if (old == 0ull) return;
            cur = old;
        }
        if ((uint32_t)cur == key) {
            atomicOr(reinterpret_cast<unsigned int*>(p) + 1, bit);
            return;
        }
        slot = (slot + 1) & msk;
    }
}

// Second hash of the pre-filter's 2-choice token table (the first is hash_u32).
__device__ __forceinline__ uint32_t hash2_u32(uint32_t key, int log_size) {
    return ((key ^ (key >> 15)) * 0x85EBCA6Bu) >> (32 - log_size);
}

// Concurrent cuckoo insertion of {key, val} (val != 0) into 2^log_nb buckets of two 64-bit slots; a slot is
// empty iff it is 0.  Every thread always holds exactly one homeless item, so concurrent evictions lose nothing.
// Returns false when the eviction chain does not end (the caller then routes the group to the exact scan).
__device__ __forceinline__ bool cuckoo_insert(unsigned long long* tab, int log_nb, uint32_t key, uint32_t val) {
    unsigned long long cur = (unsigned long long)key | ((unsigned long long)val << 32);
    uint32_t b = hash_u32(key, log_nb);
    for (int it = 0; it < 256; ++it) {
        unsigned long long* p = tab + 2 * (size_t)b;
        if (atomicCAS(p, 0ull, cur) == 0ull) return true;
        if (atomicCAS(p + 1, 0ull, cur) == 0ull) return true;
        cur = atomicExch(p + (it & 1), cur);
        const uint32_t k = (uint32_t)cur;
        const uint32_t a = hash_u32(k, log_nb), c = hash2_u32(k, log_nb);
        b = (b == a) ? c : a;
    }
    return false;
}

struct PrepParams {
    const uint4* rec;
    const uint64_t* rec_off;
    const GroupDesc* groups;
    uint8_t* arena;               // zero-filled by the host before the launch
    const int32_t* excl_ids;      // per-batch exclusion ids (adjacency rows of the queries), may be null
    int32_t* lean_fail;           // [group] set to 1 when the pre-filter's cuckoo table could not be built (may be null)
};

// one CTA per group
__global__ void __launch_bounds__(kPrepThreads) group_prep_kernel(const PrepParams P) {
    const GroupDesc& gd = P.groups[blockIdx.x];
    uint8_t* base = P.arena + gd.arena_off;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = kPrepThreads / 32;
    double* nq = reinterpret_cast<double*>(base + gd.off_nq);
    unsigned long long* t_clubs = reinterpret_cast<unsigned long long*>(base + gd.off_clubs);
    unsigned long long* t_friends = reinterpret_cast<unsigned long long*>(base + gd.off_friends);
    unsigned long long* t_adj = reinterpret_cast<unsigned long long*>(base + gd.off_adj);
    unsigned long long* t_text = reinterpret_cast<unsigned long long*>(base + gd.off_text);
    double* vals = reinterpret_cast<double*>(base + gd.off_vals);

    // phase 1: insert every element of every query of the group
    for (int L = warp; L < gd.n_lanes; L += nwarps) {
        const int row = gd.row[L];
        if (row < 0) continue;
        const uint32_t* r = reinterpret_cast<const uint32_t*>(P.rec + P.rec_off[row]);
        const uint32_t ncl = r[8], nfr = r[9], ncells = r[10], ntok = r[11];
        const uint32_t o_clubs = kHdrWords, o_friends = o_clubs + pad4(ncl), o_cells = o_friends + pad4(nfr),
                       o_keys = o_cells + 4 * ncells;
        const uint32_t bit = 1u << L;
        for (uint32_t i = lane; i < ncl; i += 32) table_insert(t_clubs, 1, gd.log_sc, r[o_clubs + i], bit);
        for (uint32_t i = lane; i < nfr; i += 32) table_insert(t_friends, 1, gd.log_sf, r[o_friends + i], bit);
        if (P.excl_ids)
            for (int i = gd.adj_begin[L] + lane; i < gd.adj_begin[L + 1]; i += 32)
                table_insert(t_adj, 1, gd.log_sa, (uint32_t)P.excl_ids[i], bit);
        for (uint32_t i = lane; i < ntok; i += 32) table_insert(t_text, 2, gd.log_st, r[o_keys + i], bit);
        for (uint32_t k = lane; k < ncells; k += 32) {
            const uint32_t t = r[o_cells + 4 * k];
            nq[t * 32 + L] = *reinterpret_cast<const double*>(r + o_cells + 4 * k + 2);
        }
    }
    __syncthreads();

    // phase 2: base[slot] = exclusive prefix sum of popc(mask) over the text table
    __shared__ uint32_t warp_sums[kPrepThreads / 32];
    __shared__ uint32_t running, n_ent;
    if (tid == 0) { running = 0; n_ent = 0; }
    __syncthreads();
    const uint32_t st = 1u << gd.log_st;
    uint32_t* t_text32 = reinterpret_cast<uint32_t*>(t_text);
    for (uint32_t c0 = 0; c0 < st; c0 += kPrepThreads) {
        const uint32_t s = c0 + tid;
        const uint32_t cnt = (s < st) ? __popc(t_text32[4 * s + 1]) : 0u;
        uint32_t inc = cnt;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(kFull, inc, o);
            if (lane >= o) inc += v;
        }
        if (lane == 31) warp_sums[warp] = inc;
        __syncthreads();
        uint32_t woff = 0;
        for (int w2 = 0; w2 < warp; ++w2) woff += warp_sums[w2];
        const uint32_t run = running;
        if (s < st) t_text32[4 * s + 2] = run + woff + inc - cnt;
        __syncthreads();
        if (tid == kPrepThreads - 1) running = run + woff + inc;
        __syncthreads();
    }

    // phase 3: per-(token, query) weights, compacted by the rank of the query inside the mask
    const uint32_t msk = st - 1u;
    for (int L = warp; L < gd.n_lanes; L += nwarps) {
        const int row = gd.row[L];
        if (row < 0) continue;
        const uint32_t* r = reinterpret_cast<const uint32_t*>(P.rec + P.rec_off[row]);
        const uint32_t ncl = r[8], nfr = r[9], ncells = r[10], ntok = r[11];
        const uint32_t o_keys = kHdrWords + pad4(ncl) + pad4(nfr) + 4 * ncells, o_w = o_keys + pad4(ntok);
        const double* w = reinterpret_cast<const double*>(r + o_w);
        for (uint32_t i = lane; i < ntok; i += 32) {
            const uint32_t key = r[o_keys + i];
            uint32_t slot = hash_u32(key, gd.log_st);
            for (;;) {
                const uint32_t k2 = t_text32[4 * slot], m2 = t_text32[4 * slot + 1];
                if (m2 != 0 && k2 == key) {
                    vals[t_text32[4 * slot + 2] + __popc(m2 & ((1u << L) - 1u))] = w[i];
                    break;
                }
                if (m2 == 0) break;   // cannot happen: the key was inserted in phase 1
                slot = (slot + 1) & msk;
            }
        }
    }

    // phase 4 (full scan only): the pre-filter's fp32 view of the text table -- a cuckoo table {key, entry + 1}, the
    // entries {mask, first value index}, and the weights divided by the query cell's L2 norm
    if (gd.exact_begin == 0) return;
    __syncthreads();
    unsigned long long* ck = reinterpret_cast<unsigned long long*>(base + gd.off_text2);
    uint2* ent = reinterpret_cast<uint2*>(base + gd.off_base);
    float* v32 = reinterpret_cast<float*>(base + gd.off_vals32);
    for (uint32_t sl = tid; sl < st; sl += kPrepThreads) {
        const uint32_t key = t_text32[4 * sl], m2 = t_text32[4 * sl + 1], b0 = t_text32[4 * sl + 2];
        if (m2 == 0u) continue;
        const uint32_t idx = atomicAdd(&n_ent, 1u);
        ent[idx] = make_uint2(m2, b0);
        const uint32_t ct = key >> kTokBits;
        uint32_t mm = m2, r = 0;
        while (mm) {
            const int L = __ffs(mm) - 1;
            mm &= mm - 1;
            const double nrm = nq[ct * 32 + L];
            v32[b0 + r] = (nrm > 0.0) ? (float)(vals[b0 + r] / nrm) : 0.0f;
            ++r;
        }
        if (!cuckoo_insert(ck, gd.log_cb, key, idx + 1u) && P.lean_fail) P.lean_fail[blockIdx.x] = 1;
    }
    // clubs and friends of the queries in one cuckoo table: {id | set << 31, query mask}
    unsigned long long* cks = reinterpret_cast<unsigned long long*>(base + gd.off_sets2);
    for (int which = 0; which < 2; ++which) {
        const unsigned long long* src = which ? t_friends : t_clubs;
        const uint32_t ns = 1u << (which ? gd.log_sf : gd.log_sc);
        for (uint32_t sl = tid; sl < ns; sl += kPrepThreads) {
            const unsigned long long e = src[sl];
            const uint32_t m2 = (uint32_t)(e >> 32);
            if (m2 == 0u) continue;
            if (!cuckoo_insert(cks, gd.log_cs, (uint32_t)e | ((uint32_t)which << 31), m2) && P.lean_fail)
                P.lean_fail[blockIdx.x] = 1;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// scoring
// ------------------------------------------------------------------------------------------------

struct ScanParams {
    const uint4* rec;
    const uint64_t* rec_off;
    const GroupDesc* groups;
    const WorkItem* items;
    const uint8_t* arena;
    const NormTables* norm;
    const double* tab_comp;      // [dim_comp][dim_comp] sigmoid(z(min/max)), index [cand][query]
    const double* tab_age;
    const int32_t* cand_rows;    // explicit candidate rows (negative = skip), or null = the row index itself
    const int32_t* group_flag;   // when set: CTAs of groups whose flag is 0 exit at once (fallback pass of the full scan)
    int group_lo, group_hi;      // CTAs of groups outside [group_lo, group_hi) exit at once (re-score split over ranks)
    int want_smem;               // >= 0: CTAs of groups whose in_smem differs exit at once (device-planned items are not sorted)
    int T;
    int K, Kp;                   // top-k and its padding to a multiple of 32
    int exclude_mode;
    uint32_t smem_arena_bytes;   // shared-memory bytes reserved for the group tables
    uint64_t* out_keys;          // EPI_TOPK: [item.out_slot][32][Kp]
    float* out_f;                // EPI_STORE: [(item.out_slot + i)][32]
    float* out_fields;           // FIELDS: [(item.out_slot + i)][32][nf]
    int nf;                      // 3 + T + 2
};

// src/recommender_similarity.cpp:18-26
__device__ __forceinline__ double pk_sigmoid(double x) {
    const double e = exp(-fabs(x));
    const double num = (x >= 0.0) ? 1.0 : e;
    return num / (1.0 + e);
}

// src/recommender_similarity.cpp:28-36 / :105-111
__device__ __forceinline__ double pk_z(const NormTables* N, int field, double s) {
    if (N->use_norm[field]) return (s - N->mean[field]) / N->sd[field];
    return 6.0 * (s - 0.5);
}

__device__ __forceinline__ uint32_t probe_set(const uint2* tab, int log_size, uint32_t key) {
    const uint32_t msk = (1u << log_size) - 1u;
    uint32_t slot = hash_u32(key, log_size);
    for (;;) {
        const uint2 e = tab[slot];
        if (e.y == 0u) return 0u;
        if (e.x == key) return e.y;
        slot = (slot + 1) & msk;
    }
}

__device__ __forceinline__ uint32_t probe_text(const uint4* tab, int log_size, uint32_t key, uint32_t* base) {
    const uint32_t msk = (1u << log_size) - 1u;
    uint32_t slot = hash_u32(key, log_size);
    for (;;) {
        const uint4 e = tab[slot];
        if (e.y == 0u) return 0u;
        if (e.x == key) { *base = e.z; return e.y; }
        slot = (slot + 1) & msk;
    }
}

// Warp-cooperative insertion of `key` into query `ql`'s descending top-k list (all lanes call it
// with the same arguments).  Rare after the first few hundred candidates, hence the simple lock.
__device__ __forceinline__ void topk_insert(uint64_t* list, int K, int Kp, uint64_t key, int* lock,
                                            volatile uint64_t* tau, int lane) {
    if (lane == 0) {
        while (atomicCAS(lock, 0, 1) != 0) __nanosleep(64);
    }
    __syncwarp();
    __threadfence_block();
    volatile uint64_t* vl = list;
    if (key > vl[K - 1]) {
        int pos = 0;
        for (int s0 = 0; s0 < Kp; s0 += 32) {
            const int i = s0 + lane;
            const bool gt = (i < K) && (vl[i] > key);
            pos += __popc(__ballot_sync(kFull, gt));
        }
        // the shortlists of the lanes of a group overlap: a key that is already in the list is not inserted twice
        const bool dup = (pos < K) && (vl[pos] == key);
        for (int s0 = Kp - 32; !dup && s0 >= 0 && s0 + 32 > pos; s0 -= 32) {
            const int i = s0 + lane;
            const bool mv = (i > pos) && (i < K);
            uint64_t v = 0;
            if (mv) v = vl[i - 1];
            __syncwarp();
            if (mv) vl[i] = v;
            __syncwarp();
        }
        if (lane == 0 && !dup) vl[pos] = key;
        __syncwarp();
        if (lane == 0) *tau = vl[K - 1];
    }
    __threadfence_block();
    __syncwarp();
    if (lane == 0) atomicExch(lock, 0);
}

template <bool SM, int EPI, bool FIELDS>
__global__ void __launch_bounds__(kScanThreads, 1) fas_scan_kernel(const ScanParams P) {
    extern __shared__ uint4 smem_u4[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(smem_u4);
    const int tid = threadIdx.x, lane = tid & 31;
    const WorkItem it = P.items[blockIdx.x];
    const GroupDesc* gd = P.groups + it.group;
    if (P.group_flag && P.group_flag[it.group] == 0) return;   // uniform for the CTA
    if (it.group < P.group_lo || it.group >= P.group_hi) return;
    if (P.want_smem >= 0 && gd->in_smem != P.want_smem) return;

    // ---- shared memory carve-up ------------------------------------------------------------
    const uint32_t arena_sm = SM ? P.smem_arena_bytes : 0u;
    NormTables* N = reinterpret_cast<NormTables*>(smem + arena_sm);
    constexpr uint32_t kNormBytes = (sizeof(NormTables) + 15u) & ~15u;
    uint64_t* lists = reinterpret_cast<uint64_t*>(smem + arena_sm + kNormBytes);       // [32][Kp]
    const int Kp = (EPI == EPI_TOPK) ? P.Kp : 0;
    uint64_t* tau = lists + 32 * Kp;                                                    // [32]
    int* locks = reinterpret_cast<int*>(tau + 32);                                      // [32]
    int* next = locks + 32;

    {   // stage the group's tables and the normalizer constants
        if (SM) {
            const uint4* src = reinterpret_cast<const uint4*>(P.arena + gd->arena_off + gd->exact_begin);
            const uint32_t n16 = (gd->arena_bytes - gd->exact_begin) >> 4;
            for (uint32_t i = tid; i < n16; i += kScanThreads) smem_u4[i] = __ldg(src + i);
        }
        const uint4* nsrc = reinterpret_cast<const uint4*>(P.norm);
        uint4* ndst = reinterpret_cast<uint4*>(N);
        for (uint32_t i = tid; i < kNormBytes / 16; i += kScanThreads) ndst[i] = __ldg(nsrc + i);
        for (int i = tid; i < 32 * Kp; i += kScanThreads) lists[i] = 0ull;
        if (tid < 32) { tau[tid] = 0ull; locks[tid] = 0; }
        if (tid == 0) *next = it.cand_begin;
    }
    __syncthreads();

    const uint8_t* tb = SM ? (smem - gd->exact_begin) : (P.arena + gd->arena_off);
    const double* nq = reinterpret_cast<const double*>(tb + gd->off_nq);
    const uint2* t_clubs = reinterpret_cast<const uint2*>(tb + gd->off_clubs);
    const uint2* t_friends = reinterpret_cast<const uint2*>(tb + gd->off_friends);
    const uint2* t_adj = reinterpret_cast<const uint2*>(tb + gd->off_adj);
    const uint4* t_text = reinterpret_cast<const uint4*>(tb + gd->off_text);
    const double* vals = reinterpret_cast<const double*>(tb + gd->off_vals);
    const int log_sc = gd->log_sc, log_sf = gd->log_sf, log_sa = gd->log_sa, log_st = gd->log_st;
    const int n_lanes = gd->n_lanes;
    const int T = P.T;

    // ---- this lane's query --------------------------------------------------------------------
    const int qrow = (lane < n_lanes) ? gd->row[lane] : -1;
    const bool q_active = qrow >= 0;
    int q_uid = 0, q_pub = -1, q_gender = -1, q_comp = 0, q_age = 0, q_r0 = -1, q_r1 = -1, q_r2 = -1, q_acnt = 0;
    uint32_t q_ncl = 0, q_nfr = 0;
    uint64_t q_mask = 0;
    double q_sqc = 0.0, q_sqf = 0.0;
    if (q_active) {
        const uint4* qr = P.rec + P.rec_off[qrow];
        const uint4 a = __ldg(qr), b = __ldg(qr + 1), c = __ldg(qr + 2), d = __ldg(qr + 3), e = __ldg(qr + 4);
        q_uid = (int)a.x; q_pub = (int)a.y; q_gender = (int)a.z; q_comp = (int)a.w;
        q_age = (int)b.x; q_r0 = (int)b.y; q_r1 = (int)b.z; q_r2 = (int)b.w;
        q_ncl = c.x; q_nfr = c.y;
        q_mask = (uint64_t)d.x | ((uint64_t)d.y << 32);
        q_acnt = (int)d.z;
        q_sqc = __hiloint2double((int)e.y, (int)e.x);
        q_sqf = __hiloint2double((int)e.w, (int)e.z);
    }
    const uint32_t lane_lt = (1u << lane) - 1u;
    const int dim_comp = N->dim_comp, dim_age = N->dim_age;
    const int total_possible = N->total_possible;
    (void)total_possible;

    // ---- candidate loop -------------------------------------------------------------------------
    // A warp claims one candidate at a time from the item's cursor in shared memory.  What bounds the kernel is the
    // latency of its dependent chains (fp64 division / exp, table probes, the record's header -> keys -> weights),
    // so it runs with as many warps as the register file allows: on the collaborative batch of the bench 512
    // threads x 111 registers take 18.2 ms, 768 x 72 14.7 ms, 1024 x 64 12.8 ms.  PK_SCAN_PREFETCH keeps two claims
    // ahead and prefetches the whole next record, one 128-byte line per lane (measured: 2 % at 512 threads, nothing
    // at 768; its ~20 registers are worth more as occupancy).
    auto claim = [&]() {
        int c = 0;
        if (lane == 0) c = atomicAdd(next, 1);
        return __shfl_sync(kFull, c, 0);
    };
    auto row_of = [&](int ci) -> int64_t {
        if (ci >= it.cand_end) return -1;
        return P.cand_rows ? (int64_t)__ldg(P.cand_rows + ci) : (int64_t)ci;
    };
#ifdef PK_SCAN_PREFETCH
    int c_cur = claim(), c_n1 = claim(), c_n2 = claim();
    int64_t r_cur = row_of(c_cur), r_n1 = row_of(c_n1), r_n2 = row_of(c_n2);
    uint64_t o_n1 = 0, e_n1 = 0, o_n2 = 0, e_n2 = 0;
    if (r_n1 >= 0) { o_n1 = __ldg(P.rec_off + r_n1); e_n1 = __ldg(P.rec_off + r_n1 + 1); }
    if (r_n2 >= 0) { o_n2 = __ldg(P.rec_off + r_n2); e_n2 = __ldg(P.rec_off + r_n2 + 1); }
    for (; c_cur < it.cand_end; ) {
        if (r_n1 >= 0 && o_n1 + 8ull * lane < e_n1) {
            const uint4* pf = P.rec + o_n1 + 8ull * lane;
            asm volatile("prefetch.global.L2 [%0];" :: "l"(pf));
        }
        const int ci = c_cur;
        const int64_t crow = r_cur;
        // rotate the claims (the loads of claim + 2 stay in flight while this candidate is scored)
        c_cur = c_n1; r_cur = r_n1;
        c_n1 = c_n2; r_n1 = r_n2; o_n1 = o_n2; e_n1 = e_n2;
        c_n2 = claim(); r_n2 = row_of(c_n2);
        if (r_n2 >= 0) { o_n2 = __ldg(P.rec_off + r_n2); e_n2 = __ldg(P.rec_off + r_n2 + 1); }
#else
    for (;;) {
        const int ci = claim();
        if (ci >= it.cand_end) break;
        const int64_t crow = row_of(ci);
#endif
        {
            if (crow < 0) continue;   // empty shortlist slot (warp-uniform)
            const uint4* cr = P.rec + P.rec_off[crow];
            const uint32_t* cw = reinterpret_cast<const uint32_t*>(cr);
            const uint4 h0 = __ldg(cr), h1 = __ldg(cr + 1), h2 = __ldg(cr + 2), h3 = __ldg(cr + 3), h4 = __ldg(cr + 4);
            const int c_uid = (int)h0.x, c_pub = (int)h0.y, c_gender = (int)h0.z, c_comp = (int)h0.w;
            const int c_age = (int)h1.x, c_r0 = (int)h1.y, c_r1 = (int)h1.z, c_r2 = (int)h1.w;
            const uint32_t c_ncl = h2.x, c_nfr = h2.y, c_ncells = h2.z, c_ntok = h2.w;
            const uint64_t c_mask = (uint64_t)h3.x | ((uint64_t)h3.y << 32);
            const int c_bcnt = (int)h3.z;
            const uint32_t o_clubs = kHdrWords, o_friends = o_clubs + pad4(c_ncl), o_cells = o_friends + pad4(c_nfr),
                           o_keys = o_cells + 4 * c_ncells, o_w = o_keys + pad4(c_ntok);
            // the first 32 token keys are wanted after the id lists: issue the load now
            uint32_t key_first = 0;
            if ((uint32_t)lane < c_ntok) key_first = __ldg(cw + o_keys + lane);

            float* fo = nullptr;
            if (FIELDS) {
                fo = P.out_fields + ((size_t)(it.out_slot + (ci - it.cand_begin)) * 32 + lane) * P.nf;
                for (int j = 0; j < P.nf; ++j) fo[j] = -1.0f;
            }

            double sum = 0.0;
            int used = 0;

            // public / gender: equality (src/recommender_similarity.cpp:38-50)
            if (q_pub >= 0 && c_pub >= 0) { sum += (q_pub == c_pub) ? N->sig1[0] : N->sig0[0]; ++used; }
            if (q_gender >= 0 && c_gender >= 0) { sum += (q_gender == c_gender) ? N->sig1[1] : N->sig0[1]; ++used; }
            // completion / age: min/max ratio (:52-68), tabulated for small values
            if (q_comp > 0 && c_comp > 0) {
                double sg;
                if (q_comp < dim_comp && c_comp < dim_comp) sg = __ldg(P.tab_comp + (size_t)c_comp * dim_comp + q_comp);
                else {
                    const int amin = min(q_comp, c_comp), amax = max(q_comp, c_comp);
                    sg = pk_sigmoid(pk_z(N, 2, (double)amin / (double)amax));
                }
                sum += sg; ++used;
            }
            if (q_age > 0 && c_age > 0) {
                double sg;
                if (q_age < dim_age && c_age < dim_age) sg = __ldg(P.tab_age + (size_t)c_age * dim_age + q_age);
                else {
                    const int amin = min(q_age, c_age), amax = max(q_age, c_age);
                    sg = pk_sigmoid(pk_z(N, 3, (double)amin / (double)amax));
                }
                sum += sg; ++used;
            }
            // region (:70-77, src/recommender.cpp:130-139)
            if (q_acnt > 0 && c_bcnt > 0) {
                const int matches = (int)(q_r0 >= 0 && c_r0 >= 0 && q_r0 == c_r0) + (int)(q_r1 >= 0 && c_r1 >= 0 && q_r1 == c_r1) +
                                    (int)(q_r2 >= 0 && c_r2 >= 0 && q_r2 == c_r2);
                sum += N->region[(q_acnt * 4 + c_bcnt) * 4 + matches];
                ++used;
                if (FIELDS) fo[0] = (float)((double)matches / (sqrt((double)q_acnt) * sqrt((double)c_bcnt)));
            }
            // The fields that need the sigma(z(s)) chain (an fp64 division, exp, another division) are QUEUED per lane --
            // {field, numerator}, in field order -- and evaluated after the candidate's last cell, entry k of every
            // lane together.  Evaluated where they arise (round 1) the chain ran once per candidate cell in which ANY
            // lane had a match, with 3-4 lanes active: 44 % of this kernel's warp instructions.  The sum then runs in
            // the reference's field order (clubs, friends, text columns ascending), sigma(z(0)) for the joint fields
            // that are not in the queue, so every double is added exactly where the reference adds it.
            // The first kQReg entries of the queue live in REGISTERS (static indices only); the rest overflow into a
            // local array that is rarely touched.  Round 2's first version kept the whole queue in local memory:
            // 624 bytes x 512 threads did not fit the L1 left beside ~200 KB of tables (local hit rate 50 %), and the
            // local loads missed half the time.  (Staging the entries of the whole warp in shared memory and evaluating
            // them 32 at a time was also built: 14.7 -> 14.1 ms at 768 threads, but +28 % instructions and 35 KB of
            // shared memory; at 1024 threads it lost to this version.)
            double qr0 = 0.0, qr1 = 0.0, qr2 = 0.0, qr3 = 0.0, qd0 = 0.0, qd1 = 0.0;
            uint32_t qfp = 0u;                       // fields of the register entries, one byte each
            double qov[2 + kMaxT - 4];
            uint8_t qfov[2 + kMaxT - 4];
            int qn = 0;
            auto q_push = [&](int f, double v) {
                if (qn == 0) qr0 = v; else if (qn == 1) qr1 = v; else if (qn == 2) qr2 = v; else if (qn == 3) qr3 = v;
                else { qov[qn - 4] = v; qfov[qn - 4] = (uint8_t)f; }
                if (qn < 4) qfp |= (uint32_t)f << (8 * qn);
                ++qn;
            };
            auto q_take = [&](int k) -> double {
                return k == 0 ? qr0 : k == 1 ? qr1 : k == 2 ? qr2 : k == 3 ? qr3 : qov[k - 4];
            };
            uint64_t matched = 0ull;
            bool has_set[2] = {false, false};
            int inter_of[2] = {0, 0};
            // clubs / friends: |{b in B : b in A}| / (sqrt|A| sqrt|B|) (:79-91, src/recommender.cpp:119-128)
            #pragma unroll
            for (int which = 0; which < 2; ++which) {
                const uint32_t c_n = which ? c_nfr : c_ncl;
                if (c_n == 0) continue;                               // warp-uniform
                const uint32_t off = which ? o_friends : o_clubs;
                const uint2* tab = which ? t_friends : t_clubs;
                const int lg = which ? log_sf : log_sc;
                int inter = 0;
                for (uint32_t e0 = 0; e0 < c_n; e0 += 32) {
                    const uint32_t i = e0 + lane;
                    uint32_t m = 0;
                    if (i < c_n) m = probe_set(tab, lg, __ldg(cw + off + i));
                    uint32_t hits = __ballot_sync(kFull, m != 0u);
                    while (hits) {
                        const int j = __ffs(hits) - 1;
                        hits &= hits - 1;
                        const uint32_t mj = __shfl_sync(kFull, m, j);
                        inter += (mj >> lane) & 1u;
                    }
                }
                const uint32_t q_n = which ? q_nfr : q_ncl;
                if (q_n > 0) {
                    has_set[which] = true;
                    inter_of[which] = inter;
                    ++used;
                    if (inter > 0) {
                        const double dd = (which ? q_sqf : q_sqc) * __hiloint2double(which ? (int)h4.w : (int)h4.y, which ? (int)h4.z : (int)h4.x);
                        if (which) qd1 = dd; else qd0 = dd;
                        q_push(5 + which, (double)inter);
                    }
                    if (FIELDS) { fo[1 + which] = 0.0f; fo[3 + T + which] = __int_as_float(inter); }
                }
            }
            // text columns (:93-114): tf-idf cosine per jointly non-empty column
            const uint64_t joint = q_mask & c_mask;
            if (c_ncells > 0) {
                // The candidate's token keys ascend over the record (column, then token), so the hits of one column
                // are consecutive: the loop runs over the HITS only and closes a column's dot product when the next
                // hit belongs to another column (round 1 walked every cell of the candidate, one dependent load of
                // the cell header each).  A lane's products are added in ascending token order, as the reference's
                // map walk adds them (src/recommender.cpp:86-100).
                const uint32_t* keys = cw + o_keys;
                const double* tokw = reinterpret_cast<const double*>(cw + o_w);
                int cur_ct = -1;
                double dot = 0.0;
                for (uint32_t cb = 0; cb < c_ntok; cb += 32) {
                    const uint32_t i = cb + lane;
                    uint32_t m = 0, base = 0, key = 0;
                    if (i < c_ntok) {
                        key = cb ? __ldg(keys + i) : key_first;
                        m = probe_text(t_text, log_st, key, &base);
                    }
                    uint32_t bits = __ballot_sync(kFull, m != 0u);
                    double wc = 0.0;
                    if (m) wc = __ldg(tokw + i);
                    while (bits) {
                        const int j = __ffs(bits) - 1;
                        bits &= bits - 1;
                        const uint32_t mj = __shfl_sync(kFull, m, j);
                        const uint32_t bj = __shfl_sync(kFull, base, j);
                        const double wj = __shfl_sync(kFull, wc, j);
                        const int ct = (int)(__shfl_sync(kFull, key, j) >> kTokBits);
                        if (ct != cur_ct) {              // warp-uniform: the previous column is complete
                            if (dot != 0.0) { q_push(kNumFixed + cur_ct, dot); matched |= 1ull << cur_ct; }
                            dot = 0.0;
                            cur_ct = ct;
                        }
                        if ((mj >> lane) & 1u) dot = __dadd_rn(dot, __dmul_rn(vals[bj + __popc(mj & lane_lt)], wj));
                    }
                }
                if (dot != 0.0) { q_push(kNumFixed + cur_ct, dot); matched |= 1ull << cur_ct; }
                used += __popcll(joint);
            }
            // evaluation pass: queue entry k of every lane of the warp together
            auto q_eval = [&](int f, double v) -> double {
                double den;
                if (f >= kNumFixed) {          // the candidate cell's norm sits at the rank of the column in its mask
                    const int ct = f - kNumFixed;
                    const uint4 cc = __ldg(cr + (o_cells >> 2) + __popcll(c_mask & ((1ull << ct) - 1ull)));
                    den = nq[ct * 32 + lane] * __hiloint2double((int)cc.w, (int)cc.z);
                } else den = (f == 5) ? qd0 : qd1;
                const float s = (den > 0.0) ? (float)(v / den) : 0.0f;
                if (FIELDS) fo[f >= kNumFixed ? 3 + (f - kNumFixed) : f - 4] = s;
                return (f >= kNumFixed && s == 1.0f) ? N->sig1[f] : pk_sigmoid(pk_z(N, f, (double)s));
            };
            if (qn > 0) qr0 = q_eval((int)(qfp & 0xffu), qr0);
            if (qn > 1) qr1 = q_eval((int)((qfp >> 8) & 0xffu), qr1);
            if (qn > 2) qr2 = q_eval((int)((qfp >> 16) & 0xffu), qr2);
            if (qn > 3) qr3 = q_eval((int)(qfp >> 24), qr3);
            for (int k = 4; k < qn; ++k) qov[k - 4] = q_eval((int)qfov[k - 4], qov[k - 4]);
            // ordered sum
            {
                int k = 0;
                if (has_set[0]) sum += (inter_of[0] > 0) ? q_take(k++) : N->sig0[5];
                if (has_set[1]) sum += (inter_of[1] > 0) ? q_take(k++) : N->sig0[6];
                uint64_t jm = (c_ncells > 0) ? joint : 0ull;
                while (jm) {
                    const int ct = __ffsll((long long)jm) - 1;
                    jm &= jm - 1;
                    if ((matched >> ct) & 1ull) sum += q_take(k++);
                    else {
                        sum += N->sig0[kNumFixed + ct];
                        if (FIELDS) fo[3 + ct] = 0.0f;
                    }
                }
            }

            // S, F, harmonic mean (:116-123)
            float fas = 0.0f;
            if (used > 0) {
                const double S = sum / (double)used;
                const double F = N->ftab[used];
                fas = (float)((2.0 * S * F) / (S + F));
            }

            if (EPI == EPI_STORE) {
                P.out_f[(size_t)(it.out_slot + (ci - it.cand_begin)) * 32 + lane] = q_active ? fas : -1.0f;
            } else {
                bool ok = q_active;
                if (P.exclude_mode >= PK_EXCLUDE_SELF && c_uid == q_uid) ok = false;
                if (P.exclude_mode >= PK_EXCLUDE_SELF_AND_ADJ) {
                    const uint32_t mx = probe_set(t_adj, log_sa, (uint32_t)c_uid);   // warp-uniform probe
                    if ((mx >> lane) & 1u) ok = false;
                }
                const uint64_t key = make_key(__float_as_uint(fas), c_uid);
                const uint64_t tl = reinterpret_cast<volatile uint64_t*>(tau)[lane];
                uint32_t pass = __ballot_sync(kFull, ok && key > tl);
                while (pass) {
                    const int j = __ffs(pass) - 1;
                    pass &= pass - 1;
                    const uint64_t kj = __shfl_sync(kFull, key, j);
                    topk_insert(lists + j * Kp, P.K, Kp, kj, locks + j, reinterpret_cast<volatile uint64_t*>(tau) + j, lane);
                }
            }
        }
    }

    if (EPI == EPI_TOPK) {
        __syncthreads();
        uint64_t* dst = P.out_keys + (size_t)it.out_slot * 32 * Kp;
        for (int i = tid; i < 32 * Kp; i += kScanThreads) dst[i] = lists[i];
    }
}

// ------------------------------------------------------------------------------------------------
// merge / select: per query, the K best DISTINCT keys of a set of key segments (sorted or not)
// ------------------------------------------------------------------------------------------------

struct MergeParams {
    const uint64_t* keys;
    // addressing mode A (scan output): list j of query q = keys + ((slot0[q] + j) * 32 + lane[q]) * stride_k, K entries
    // addressing mode B (gathered):    list j of query q = keys + ((size_t)j * nq + q) * stride_k, K entries
    // addressing mode C (slice):       one unsorted segment keys[seg_begin[q] .. seg_begin[q + 1])
    const int32_t* slot0;       // mode A: first out_slot of the query's group
    const int32_t* n_lists_q;   // mode A: number of lists of each query
    const int32_t* lane;        // mode A
    const int32_t* seg_begin;   // mode C
    const int32_t* seg_cnt;     // mode C: length of each segment (null = up to the next segment's begin)
    int n_lists;                // mode B
    int nq, K, stride_k;
    int cap;                    // sort buffer capacity in keys (power of two, >= 2 K and >= kMergeThreads)
    uint64_t* out_keys;         // [nq][K], 0-padded
    // mode A, two-stage full scan: queries whose flag is set take their lists from the fallback pass instead
    const int32_t* q_flag;
    const uint64_t* alt_keys;
    const int32_t* alt_slot0;
    const int32_t* alt_n_lists_q;
    int alt_stride_k;
    // pre-filter shortlists: a query whose group's fp32 tables could not be built on THIS rank gets the poison key
    // ~0 in front of its list; it survives every later merge (it is the largest key), so every rank sees it and
    // flags the query in shortlist_check_kernel -- the ranks' cuckoo builds need not fail alike
    const int32_t* poison_group_fail;   // [group] (null = no poisoning)
    const int32_t* q_group;
};

constexpr int kMergeThreads = 256;
constexpr uint64_t kPoisonKey = ~0ull;

// One CTA per query.  The candidates are pulled through a shared-memory buffer in rounds: [running top-K | next
// keys], bitonic sort (descending), duplicates dropped (the lists of a re-scored shortlist overlap: equal keys
// are adjacent after the sort), first K kept.  ~log^2(cap)/2 compare-exchange steps per round instead of K
// block-wide arg-max rounds with two barriers each.
__global__ void __launch_bounds__(kMergeThreads) merge_topk_kernel(const MergeParams P) {
    extern __shared__ uint64_t merge_sm[];
    uint64_t* buf = merge_sm;               // [cap]
    uint64_t* res = merge_sm + P.cap;       // [K]
    __shared__ int wsum[kMergeThreads / 32];
    const int q = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // a flagged query (two-stage full scan) always takes its lists from the fallback pass, in mode-A addressing
    const bool alt = P.q_flag != nullptr && P.alt_keys != nullptr && P.q_flag[q] != 0;
    const bool modeC = !alt && P.seg_begin != nullptr;
    const bool modeA = alt || (!modeC && P.slot0 != nullptr);
    const uint64_t* kbase = alt ? P.alt_keys : P.keys;
    const int stride = alt ? P.alt_stride_k : P.stride_k;
    const int K = P.K;
    int64_t total;
    const uint64_t* seg = nullptr;
    size_t list0 = 0;
    if (modeC) {
        const int b = P.seg_begin[q];
        seg = P.keys + b;
        total = P.seg_cnt ? P.seg_cnt[q] : P.seg_begin[q + 1] - b;
    } else if (modeA) {
        const int nl = alt ? P.alt_n_lists_q[q] : P.n_lists_q[q];
        list0 = ((size_t)(alt ? P.alt_slot0 : P.slot0)[q] * 32 + P.lane[q]) * stride;
        total = (int64_t)nl * K;
    } else {
        total = (int64_t)P.n_lists * K;
    }
    auto fetch = [&](int64_t idx) -> uint64_t {
        if (modeC) return seg[idx];
        const int64_t j = idx / K;
        const int i = (int)(idx - j * K);
        if (modeA) return kbase[list0 + (size_t)j * 32 * stride + i];
        return P.keys[((size_t)j * P.nq + q) * P.stride_k + i];
    };

    int have = 0;
    for (int64_t base = 0; base < total;) {
        const int64_t left_ = total - base, room_ = (int64_t)(P.cap - have);
        const int take = (int)(left_ < room_ ? left_ : room_);
        for (int i = tid; i < take; i += kMergeThreads) buf[have + i] = fetch(base + i);
        const int n = have + take;
        int n2 = kMergeThreads;
        while (n2 < n) n2 <<= 1;
        for (int i = n + tid; i < n2; i += kMergeThreads) buf[i] = 0ull;
        __syncthreads();
        // bitonic sort, descending.  Steps with j <= 16 exchange inside a warp's own 32 consecutive pair slots.
        for (int k = 2; k <= n2; k <<= 1) {
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int t = tid; t < (n2 >> 1); t += kMergeThreads) {
                    const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), l = i | j;
                    const bool desc = (i & k) == 0;
                    const uint64_t a = buf[i], b = buf[l];
                    if ((a < b) == desc) { buf[i] = b; buf[l] = a; }
                }
                const int j_next = (j > 1) ? (j >> 1) : k;
                if (j_next > 16) __syncthreads(); else __syncwarp();
            }
        }
        __syncthreads();
        // drop zeros and duplicates, keep the first K (order-preserving compaction)
        const int per = n2 / kMergeThreads, i0 = tid * per;
        int cnt = 0;
        for (int i = i0; i < i0 + per; ++i) {
            const uint64_t v = buf[i];
            cnt += (v != 0ull && (i == 0 || buf[i - 1] != v)) ? 1 : 0;
        }
        int inc = cnt;
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(kFull, inc, o);
            if (lane >= o) inc += v;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        int pos = inc - cnt, tot = 0;
        for (int w2 = 0; w2 < kMergeThreads / 32; ++w2) { if (w2 < warp) pos += wsum[w2]; tot += wsum[w2]; }
        for (int i = i0; i < i0 + per && pos < K; ++i) {
            const uint64_t v = buf[i];
            if (v != 0ull && (i == 0 || buf[i - 1] != v)) res[pos++] = v;
        }
        have = min(tot, K);
        base += take;
        __syncthreads();
        if (base < total)
            for (int i = tid; i < have; i += kMergeThreads) buf[i] = res[i];
        __syncthreads();
    }
    const bool poison = P.poison_group_fail != nullptr && P.q_group[q] >= 0 && P.poison_group_fail[P.q_group[q]] != 0;
    uint64_t* out = P.out_keys + (size_t)q * K;
    for (int i = tid; i < K; i += kMergeThreads) {
        uint64_t v = 0ull;
        if (poison) v = (i == 0) ? kPoisonKey : ((i - 1 < have) ? res[i - 1] : 0ull);
        else if (i < have) v = res[i];
        out[i] = v;
    }
}

// collaborative reduce (reference src/recommender_graph.cpp:167-180,213): for every candidate of a
// user, score = sum over the user's friends IN ADJACENCY ORDER of (double)w_f * (double)FAS(f, cand),
// rounded to float once; emitted as sort keys.  `fas` is the output of the store-mode scan: w_f of
// distinct friend `sl` sits at user_w_off + sl*32 (1-lane group {user} x friends), FAS(f, cand) at
// user_mat_off + ((sl/32)*n_cand + cand)*32 + sl%32 (one [cand][32] block per 32 distinct friends).
struct CollabReduceParams {
    const float* fas;
    const int32_t* user_cand_begin;   // [nu+1] candidate slice of each user (global candidate index)
    const int32_t* user_fr_begin;     // [nu+1] slice of fr_slot
    const int32_t* fr_slot;           // per (user, friend in adjacency order): distinct-friend slot, -1 = no profile
    const int64_t* user_w_off;        // [nu]
    const int64_t* user_mat_off;      // [nu]
    const int32_t* cand_uid;          // per global candidate index
    uint64_t* out_keys;               // per global candidate index
    int nu;
};

__global__ void collab_reduce_kernel(const CollabReduceParams P) {
    const int u = blockIdx.y;
    const int cb = P.user_cand_begin[u], ce = P.user_cand_begin[u + 1];
    const int fb = P.user_fr_begin[u], fe = P.user_fr_begin[u + 1];
    const int64_t ncand = ce - cb;
    const float* wv = P.fas + P.user_w_off[u];
    const float* mat = P.fas + P.user_mat_off[u];
    for (int c = cb + blockIdx.x * blockDim.x + threadIdx.x; c < ce; c += gridDim.x * blockDim.x) {
        double score = 0.0;
        for (int f = fb; f < fe; ++f) {
            const int sl = P.fr_slot[f];
            if (sl < 0) continue;
            const float w = wv[(int64_t)sl * 32];
            const float sfc = mat[((int64_t)(sl >> 5) * ncand + (c - cb)) * 32 + (sl & 31)];
            score = __dadd_rn(score, __dmul_rn((double)w, (double)sfc));
        }
        P.out_keys[c] = make_key(__float_as_uint((float)score), P.cand_uid[c]);
    }
}

}  // namespace pk
