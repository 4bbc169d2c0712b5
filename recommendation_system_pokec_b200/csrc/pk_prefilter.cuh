// pk_prefilter.cuh -- the fp32 pre-filter of the full scan.
//
// The full scan asks for the top-k of 1.6 M exact FAS values per query, but only the few candidates near
// the k-th score need the reference's fp64 arithmetic.  fas_prefilter_kernel streams the LEAN records
// (pk_layout.h) and computes, for every (query, candidate) pair, an fp32 approximation of
// Recommender::profile_similarity (reference src/recommender_similarity.cpp:10-124) whose distance from the
// exact value is bounded by LeanNorm::margin.  It keeps the Ks > k best approximate scores per query.  If
// the Ks-th approximate score lies more than 2*margin below the k-th, the exact top-k is a subset of that
// shortlist (shortlist_check_kernel verifies this per query); the shortlist is then re-scored by the exact
// kernel (fas_scan_kernel in list mode).  Queries that fail the check are re-run by the exact full scan.
// The result is therefore always the exact kernel's; the pre-filter only decides which pairs it sees.
#pragma once
#include "pk_scan.cuh"

namespace pk {

#ifdef PK_PRE_THREADS
constexpr int kPreThreads = PK_PRE_THREADS;      // experiment builds (build.build_variant)
#else
constexpr int kPreThreads = 896;                 // 28 warps x 72 registers: measured 1.2 % faster than 1024 x 64 (spills), 768 x 77 is 8 % slower
#endif
constexpr int kPreWarps = kPreThreads / 32;

// one warp per row: lean record from the packed (exact) record
__global__ void lean_pack_kernel(int n, const uint4* __restrict__ rec, const uint64_t* __restrict__ rec_off,
                                 const uint64_t* __restrict__ lrec_off, uint4* __restrict__ lrec, uint2* __restrict__ lmask) {
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= n) return;
    const uint32_t* r = reinterpret_cast<const uint32_t*>(rec + rec_off[row]);
    uint32_t* out = reinterpret_cast<uint32_t*>(lrec + lrec_off[row]);
    const uint32_t ncl = r[8], nfr = r[9], ncells = r[10], ntok = r[11];
    const uint32_t o_clubs = kHdrWords, o_friends = o_clubs + pad4(ncl), o_cells = o_friends + pad4(nfr),
                   o_keys = o_cells + 4 * ncells, o_w = o_keys + pad4(ntok);
    const uint32_t l_clubs = kLeanHdrWords, l_friends = l_clubs + pad4(ncl), l_keys = l_friends + pad4(nfr),
                   l_w = l_keys + pad4(ntok);
    if (lane < 8) out[lane] = r[lane];
    if (lane == 8) out[8] = ncl;
    if (lane == 9) out[9] = nfr;
    if (lane == 10) out[10] = ntok;
    if (lane == 11) out[11] = r[14];
    if (lane == 12) out[12] = r[12];
    if (lane == 16) lmask[row] = make_uint2(r[12], r[13]);     // the column masks again, 8 bytes per row (level-1 test)
    if (lane == 13) out[13] = r[13];
    if (lane == 14) out[14] = __float_as_uint(ncl ? (float)(1.0 / sqrt((double)ncl)) : 0.0f);
    if (lane == 15) out[15] = __float_as_uint(nfr ? (float)(1.0 / sqrt((double)nfr)) : 0.0f);
    for (uint32_t i = lane; i < pad4(ncl); i += 32) out[l_clubs + i] = r[o_clubs + i];
    for (uint32_t i = lane; i < pad4(nfr); i += 32) out[l_friends + i] = r[o_friends + i];
    for (uint32_t i = lane; i < pad4(ntok); i += 32) {
        out[l_keys + i] = (i < ntok) ? r[o_keys + i] : 0xffffffffu;
        out[l_w + i] = 0u;
    }
    __syncwarp();
    const double* w = reinterpret_cast<const double*>(r + o_w);
    uint32_t begin = 0;
    for (uint32_t c = 0; c < ncells; ++c) {
        const uint32_t end = r[o_cells + 4 * c + 1];
        const double nrm = *reinterpret_cast<const double*>(r + o_cells + 4 * c + 2);
        for (uint32_t i = begin + lane; i < end; i += 32)
            out[l_w + i] = __float_as_uint((nrm > 0.0) ? (float)(w[i] / nrm) : 0.0f);
        begin = end;
    }
}

struct PreParams {
    const uint4* lrec;
    const uint64_t* lrec_off;
    const uint2* lmask;          // [row] 64-bit mask of non-empty text columns (8 bytes per candidate: the level-1 test)
    // Scan order.  The first n_list positions of a group's cursor are the shard's HEAVIEST rows (row_list, most tokens
    // first-class candidates for every top-k), the following positions the shard in row order minus those rows
    // (skip[row] != 0).  The thresholds theta are then close to their final values after a few hundred candidates
    // instead of tens of thousands -- the warm-up is a fixed cost per launch, i.e. what an 8-GPU shard pays 8x as often.
    const int32_t* row_list;
    const uint8_t* skip;
    int n_list;
    const GroupDesc* groups;
    const WorkItem* items;
    const uint8_t* arena;
    const LeanNorm* ln;
    const float* tab_comp;       // [dim][dim] float copies of the completion / age tables, index [cand][query]
    const float* tab_age;
    int Ks;                      // shortlist length (multiple of 32)
    int exclude_mode;
    uint32_t smem_arena_bytes;
    // TOPK: dynamic scheduling.  The grid is one persistent CTA per SM.  A CTA picks a group (first its own, later
    // the one with the most candidates left), takes one of the group's `max_inst` shortlist instances, claims chunks
    // of candidates from the group's cursor until the shard is exhausted, writes its lists to
    // out_keys[group_index * max_inst + instance] and picks again.  Far fewer list instances than a static slicing
    // that fills the GPU (each instance pays ~Ks (1 + ln(n / Ks)) insertions), and no wave tail.
    int* dyn_cursor;             // [n_dyn] next candidate offset inside the shard (zeroed per launch)
    int* dyn_ninst;              // [n_dyn] instances handed out (zeroed per launch)
    const int32_t* dyn_groups;   // [n_dyn] group index of each schedulable group
    int n_dyn, max_inst;
    // Header-only pruning.  theta = (k-th best approximate score so far) - 2 margin - slack: a candidate whose UPPER
    // BOUND (exact scalar fields, every jointly filled text column / set field at its largest sigma) lies below
    // theta for all 32 queries is skipped after its 64-byte header -- it cannot reach the exact top-k, so the
    // shortlist does not need it (the proof of shortlist_check_kernel only needs the candidates >= a_k - 2 margin).
    // dyn_theta[group][32] carries the best theta any instance of the group has reached: a 64-bit word
    // {theta_tag, float bits}, raised with atomicMax (the array is zeroed per execution, so the tag is constant in it).
    // peer_theta[r] are the same arrays of the OTHER ranks of a sharded scan, mapped through NVLink peer memory
    // (cudaIpc): warp 0 of every CTA polls one peer per chunk and folds what it finds into its own array.  Every
    // rank's k-th best score is a lower bound of the merged k-th best, so a peer's theta is valid on this shard too --
    // if it belongs to the same execution of the same plan, which is what the tag (plan fingerprint + execution
    // counter) says; anything else is ignored.
    unsigned long long* dyn_theta;
    const unsigned long long* peer_theta[7];
    int n_peers;
    unsigned int theta_tag;
    // The k-th best of ONE shard says little about the k-th best of all shards (of the global top-50 a shard of eight
    // holds ~6).  So every rank also publishes, per query, the score at its local rank k_frac = ceil(k / ranks)
    // (dyn_sfrac, same tagging; the peers' arrays sit at the same offset of their mapped allocations): the ranks
    // together hold at least ranks * k_frac >= k candidates scoring at least the MINIMUM of those values, which
    // therefore bounds the merged k-th best from below -- a far tighter threshold than any single rank's own k-th best.
    unsigned long long* dyn_sfrac;
    const unsigned long long* peer_sfrac[7];
    int k_frac;
    int K;                       // the caller's k (theta watches list position K - 1)
    float theta_drop;            // 2 margin + slack
    int cand_begin, cand_end;    // the shard's row range
    uint64_t* out_keys;          // TOPK: [item.out_slot][32][Ks], key = (approx score bits, candidate ROW)
    float* out_f;                // STORE (tests): [(item.out_slot + i)][32] approximate scores
};

__device__ __forceinline__ float lean_sigma(float4 c, float s) {
    // 1 / (1 + 2^(a s + b)); ex2 saturates to +inf / 0, the reciprocal to 0 / 1
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fmaf(c.x, s, c.y)));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
    return r;
}

// Level-1 bound of the header-only pruning: which counts j of jointly filled text columns could still reach theta
// for this query, whatever the candidate's scalar fields are.  frow = {n, the largest sigmas of the query's own
// n fixed / set fields, DESCENDING}: a candidate shares some subset of them -- for every subset size p the best
// case is the top-p prefix -- plus j text columns at ub_text.  Writes the 64-bit mask of admissible j into q4->z/w
// and the smallest bound among the admissible counts into *bnd (the mask is unchanged while theta stays below it).
__device__ __forceinline__ float l1_bound(int j, const float* frow, int nf, float ub_text, float inv_total) {
    float pre = 0.f, best = 0.f;
    for (int p = 0; p <= nf; ++p) {
        const float sm = fmaf((float)j, ub_text, pre), u = (float)(p + j);
        if (u > 0.f) {
            const float S = sm / u, F = u * inv_total;
            best = fmaxf(best, 2.0f * S * F / (S + F));
        }
        if (p < nf) pre += frow[1 + p];
    }
    return best * (1.0f + 2e-6f) + 2e-6f;
}
__device__ __noinline__ void l1_update(float th, const float* frow, float ub_text, float inv_total, volatile uint4* q4,
                                       volatile float* bnd) {
    const int nf = __float_as_int(frow[0]);
    uint32_t lo = 0u, hi = 0u;
    float nb = 3.0e38f;
    for (int j = 0; j <= kMaxT; ++j) {
        const float best = l1_bound(j, frow, nf, ub_text, inv_total);
        if (best >= th) {
            const int bit = min(j, 63);                 // j = 64 (T = 64, every column joint) shares bit 63
            if (bit < 32) lo |= 1u << bit; else hi |= 1u << (bit - 32);
            nb = fminf(nb, best);
        }
    }
    q4->z = lo; q4->w = hi;
    *bnd = nb;
}

// the same update by a whole warp (lane l evaluates the counts l and l + 32): used after an insertion moved a threshold
__device__ __forceinline__ void l1_update_warp(float th, const float* frow, float ub_text, float inv_total, volatile uint4* q4,
                                               volatile float* bnd, int lane) {
    const int nf = __float_as_int(frow[0]);
    const float b0 = l1_bound(lane, frow, nf, ub_text, inv_total);
    float b1 = l1_bound(lane + 32, frow, nf, ub_text, inv_total);
    if (lane == 31) b1 = fmaxf(b1, l1_bound(64, frow, nf, ub_text, inv_total));      // j = 64 shares bit 63
    const uint32_t lo = __ballot_sync(kFull, b0 >= th), hi = __ballot_sync(kFull, b1 >= th);
    float nb = fminf(b0 >= th ? b0 : 3.0e38f, b1 >= th ? b1 : 3.0e38f);
    for (int o = 16; o; o >>= 1) nb = fminf(nb, __shfl_xor_sync(kFull, nb, o));
    if (lane == 0) { q4->z = lo; q4->w = hi; *bnd = nb; }
}

// Warp-cooperative insertion of `key` into one query's descending shortlist (all lanes call it with the same
// arguments; keys are unique -- one per candidate row).
__device__ __forceinline__ void shortlist_insert(uint64_t* list, int K, uint64_t key, int* lock, volatile uint64_t* tau,
                                                 int lane, int k_user, float theta_drop, volatile float* theta,
                                                 unsigned long long* g_theta, unsigned int tag, unsigned long long* g_sfrac, int k_frac,
                                                 const float* frow, float ub_text, float inv_total,
                                                 volatile uint4* q4, volatile float* bnd) {
    if (lane == 0) {
        while (atomicCAS(lock, 0, 1) != 0) __nanosleep(32);
    }
    __syncwarp();
    __threadfence_block();
    volatile uint64_t* vl = list;
    float th_moved = -1.f;                 // lane 0: the new pruning threshold, when the insertion moved the k-th best
    if (key > vl[K - 1]) {
        int pos = 0;
        for (int s0 = 0; s0 < K; s0 += 32) pos += __popc(__ballot_sync(kFull, vl[s0 + lane] > key));
        for (int s0 = K - 32; s0 >= 0 && s0 + 32 > pos; s0 -= 32) {
            const int i = s0 + lane;
            const bool mv = i > pos;
            uint64_t v = 0;
            if (mv) v = vl[i - 1];
            __syncwarp();
            if (mv) vl[i] = v;
            __syncwarp();
        }
        if (lane == 0) {
            vl[pos] = key;
            *tau = vl[K - 1];
            if (g_sfrac) {                           // sharded scan: the score at local rank k_frac, for the other ranks
                const uint64_t kf = vl[k_frac - 1];
                if (kf != 0ull && pos < k_frac) atomicMax(g_sfrac, ((unsigned long long)tag << 32) | key_score_bits(kf));
            }
            const uint64_t kk = vl[k_user - 1];
            if (kk != 0ull && pos < k_user) {      // the k-th best moved: raise this query's pruning threshold
                const float th = fmaxf(__uint_as_float(key_score_bits(kk)) - theta_drop, 0.f);
                if (th > *theta) {
                    *theta = th;
                    if (g_theta) atomicMax(g_theta, ((unsigned long long)tag << 32) | __float_as_uint(th));
                    th_moved = th;
                }
            }
        }
    }
    __threadfence_block();
    __syncwarp();
    if (lane == 0) atomicExch(lock, 0);
    // the level-1 mask follows outside the lock (a stale mask only admits more candidates)
    // (lane 0 decides: *bnd may be rewritten by another warp between two lanes' reads, and the update uses full-warp votes)
    th_moved = (lane == 0 && th_moved > *bnd) ? th_moved : -1.f;
    th_moved = __shfl_sync(kFull, th_moved, 0);
    if (th_moved >= 0.f) l1_update_warp(th_moved, frow, ub_text, inv_total, q4, bnd, lane);
}

// 32 x 32 bit transpose across the warp: bit j of the result in lane l = bit l of x in lane j.
// Butterfly over s = 16 .. 1 (swap the off-diagonal s x s sub-blocks of every 2s x 2s block) carried out in a
// rotated frame: lane l holds its row rotated left by l.  In that frame the partner's row needs no shift at all
// (upper rows would shift up by s, lower rows down by s; the relative rotation l - (l ^ s) cancels both), and the
// columns a lane keeps are rotl(lo_s, l mod s) for upper and lower rows alike -- so a step is one SHFL and one LOP3.
__device__ __forceinline__ uint32_t transpose32(uint32_t x, int lane) {
    x = __funnelshift_l(x, x, lane);
    #pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        const uint32_t lo = (s == 16) ? 0x0000ffffu : (s == 8) ? 0x00ff00ffu : (s == 4) ? 0x0f0f0f0fu : (s == 2) ? 0x33333333u : 0x55555555u;
        // (the four lane-dependent masks are rebuilt per call: fetching them from shared memory measured slower)
        const uint32_t keep = __funnelshift_l(lo, lo, lane & (s - 1));
        const uint32_t y = __shfl_xor_sync(kFull, x, s);
        x = (x & keep) | (y & ~keep);
    }
    return __funnelshift_r(x, x, lane);
}

template <bool STORE>
__global__ void __launch_bounds__(kPreThreads, 1) fas_prefilter_kernel(const PreParams P) {
    extern __shared__ uint4 smem_u4[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(smem_u4);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t arena_sm = P.smem_arena_bytes;
    LeanNorm* N = reinterpret_cast<LeanNorm*>(smem + arena_sm);
    constexpr uint32_t kNormBytes = (sizeof(LeanNorm) + 15u) & ~15u;
    const int Ks = STORE ? 0 : P.Ks;
    uint64_t* lists = reinterpret_cast<uint64_t*>(smem + arena_sm + kNormBytes);   // [32][Ks]
    uint64_t* tau = lists + 32 * Ks;
    int* locks = reinterpret_cast<int*>(tau + 32);
    float* theta = reinterpret_cast<float*>(locks + 32);                            // [32] pruning thresholds
    uint4* l1q = reinterpret_cast<uint4*>(theta + 32);                              // [32] {query mask lo, hi, jmin, -}
    float* l1f = reinterpret_cast<float*>(l1q + 32);                                // [32][8] {n, largest sigmas descending}
    float* l1b = l1f + 32 * 8;                                                      // [32] theta up to which l1q[].z/w hold
    float* pcache = l1b + 32;                                                       // [7][32] the peers' rank-k_frac scores, as last polled
    int* next = reinterpret_cast<int*>(pcache + 7 * 32);
    uint4* stage = reinterpret_cast<uint4*>(next + 4) + warp * 32;                  // per-warp hit staging
    int* pick = next + 1;                                                           // {group index, instance}

    for (int round = 0;; ++round) {
    WorkItem it;
    if (STORE) {
        if (round > 0) break;
        it = P.items[blockIdx.x];
    } else {
        // ---- pick a group and one of its shortlist instances (PreParams: dynamic scheduling) ----
        if (warp == 0) {
            const int n_cand = P.n_list + (P.cand_end - P.cand_begin);      // cursor positions
            int g = -1, inst = -1;
            for (int attempt = 0; attempt < 8 && g < 0; ++attempt) {
                int best = -1, best_rem = 0;
                if (round == 0 && attempt == 0) {
                    best = (int)(blockIdx.x % (unsigned)P.n_dyn);     // the first pick spreads the CTAs over the groups
                    best_rem = n_cand;
                } else {
                    // the group with the most candidates left that still has a free instance; joining for less than
                    // a few thousand candidates would cost more (tables, list fill, merge) than it saves
                    // (a group nobody has started yet is always taken, however small)
                    for (int i = lane; i < P.n_dyn; i += 32) {
                        const int rem = n_cand - *reinterpret_cast<volatile int*>(P.dyn_cursor + i);
                        const int ni = *reinterpret_cast<volatile int*>(P.dyn_ninst + i);
                        const bool worth = rem > 4096 || (ni == 0 && rem > 0);
                        if (worth && rem > best_rem && ni < P.max_inst) { best_rem = rem; best = i; }
                    }
                    for (int o = 16; o >= 1; o >>= 1) {
                        const int r2 = __shfl_xor_sync(kFull, best_rem, o), b2 = __shfl_xor_sync(kFull, best, o);
                        if (r2 > best_rem || (r2 == best_rem && b2 > best)) { best_rem = r2; best = b2; }
                    }
                }
                if (best < 0) break;
                int i = 0;
                if (lane == 0) i = atomicAdd(P.dyn_ninst + best, 1);
                i = __shfl_sync(kFull, i, 0);
                if (i < P.max_inst) { g = best; inst = i; }
            }
            if (lane == 0) { pick[0] = g; pick[1] = inst; }
        }
        __syncthreads();
        const int gi = pick[0], inst = pick[1];
        if (gi < 0) break;
        it.group = P.dyn_groups[gi];
        it.cand_begin = 0; it.cand_end = P.n_list + (P.cand_end - P.cand_begin);        // cursor positions, see PreParams
        it.out_slot = gi * P.max_inst + inst;
    }
    const GroupDesc* gd = P.groups + it.group;
    int* const cursor = STORE ? next : P.dyn_cursor + pick[0];
    const int cursor_base = 0;

    {
        const uint4* src = reinterpret_cast<const uint4*>(P.arena + gd->arena_off);
        const uint32_t n16 = gd->lean_end >> 4;
        for (uint32_t i = tid; i < n16; i += kPreThreads) smem_u4[i] = __ldg(src + i);
        const uint4* nsrc = reinterpret_cast<const uint4*>(P.ln);
        uint4* ndst = reinterpret_cast<uint4*>(N);
        for (uint32_t i = tid; i < kNormBytes / 16; i += kPreThreads) ndst[i] = __ldg(nsrc + i);
        for (int i = tid; i < 32 * Ks; i += kPreThreads) lists[i] = 0ull;
        if (tid < 32) { tau[tid] = 0ull; locks[tid] = 0; theta[tid] = 0.f; }
        if (tid < 7 * 32) pcache[tid] = 0.f;
        if (tid == 0) *next = it.cand_begin;
    }
    __syncthreads();

    const uint4* t_ck = reinterpret_cast<const uint4*>(smem + gd->off_text2);      // cuckoo buckets {key, entry+1} x 2
    const uint2* t_ent = reinterpret_cast<const uint2*>(smem + gd->off_base);      // entries {query mask, first value}
    const float* vals = reinterpret_cast<const float*>(smem + gd->off_vals32);
    const uint4* t_sets = reinterpret_cast<const uint4*>(smem + gd->off_sets2);     // clubs + friends, cuckoo buckets
    const uint2* t_adj = reinterpret_cast<const uint2*>(smem + gd->off_adj);
    const int log_cs = gd->log_cs, log_sa = gd->log_sa, log_cb = gd->log_cb;

    // ---- this lane's query ----------------------------------------------------------------------
    const int qrow = (lane < gd->n_lanes) ? gd->row[lane] : -1;
    const bool q_active = qrow >= 0;
    int q_uid = 0, q_pub = -1, q_gender = -1, q_comp = 0, q_age = 0, q_r0 = -1, q_r1 = -1, q_r2 = -1, q_acnt = 0;
    uint32_t q_ncl = 0, q_nfr = 0, q_m0 = 0, q_m1 = 0;
    float q_rsc = 0.f, q_rsf = 0.f;
    if (q_active) {
        const uint4* qr = P.lrec + P.lrec_off[qrow];
        const uint4 a = __ldg(qr), b = __ldg(qr + 1), c = __ldg(qr + 2), d = __ldg(qr + 3);
        q_uid = (int)a.x; q_pub = (int)a.y; q_gender = (int)a.z; q_comp = (int)a.w;
        q_age = (int)b.x; q_r0 = (int)b.y; q_r1 = (int)b.z; q_r2 = (int)b.w;
        q_ncl = c.x; q_nfr = c.y; q_acnt = (int)c.w;
        q_m0 = d.x; q_m1 = d.y; q_rsc = __uint_as_float(d.z); q_rsf = __uint_as_float(d.w);
    }
    const uint32_t lanebit = 1u << lane, lane_lt = lanebit - 1u;
    const int dim_comp = N->dim_comp, dim_age = N->dim_age;
    const float inv_total = N->inv_total;
    const float ub_text = N->ub_text, ub_clubs = N->ub_clubs, ub_friends = N->ub_friends;
    if (!STORE && warp == 0) {
        // level-1 tables of the group: the query's column mask, and the largest sigma of each fixed / set field the
        // query has (a candidate can only share those), sorted descending for jmin_of
        float fm[7];
        int nfm = 0;
        if (q_active) {
            const float e = 2e-6f;
            if (q_pub >= 0) fm[nfm++] = fmaxf(N->fld[0].z, N->fld[0].w) + e;
            if (q_gender >= 0) fm[nfm++] = fmaxf(N->fld[1].z, N->fld[1].w) + e;
            if (q_comp > 0) fm[nfm++] = fmaxf(N->fld[2].z, N->fld[2].w) + e;
            if (q_age > 0) fm[nfm++] = fmaxf(N->fld[3].z, N->fld[3].w) + e;
            if (q_acnt > 0) fm[nfm++] = fmaxf(N->fld[4].z, N->fld[4].w) + e;
            if (q_ncl > 0) fm[nfm++] = ub_clubs;
            if (q_nfr > 0) fm[nfm++] = ub_friends;
            for (int a = 1; a < nfm; ++a)
                for (int b = a; b > 0 && fm[b] > fm[b - 1]; --b) { const float t = fm[b]; fm[b] = fm[b - 1]; fm[b - 1] = t; }
        }
        l1f[lane * 8] = __int_as_float(nfm);
        for (int a = 0; a < 7; ++a) l1f[lane * 8 + 1 + a] = (a < nfm) ? fm[a] : 0.f;
        l1q[lane] = q_active ? make_uint4(q_m0, q_m1, 0xffffffffu, 0xffffffffu) : make_uint4(0u, 0u, 0u, 0u);
        l1b[lane] = q_active ? 0.f : 3.0e38f;        // theta = 0: every count admissible
    }
    __syncthreads();
    unsigned long long* const g_theta = STORE ? nullptr : P.dyn_theta + (size_t)pick[0] * 32 + lane;
    unsigned long long* const g_sfrac = (STORE || P.n_peers == 0) ? nullptr : P.dyn_sfrac + (size_t)pick[0] * 32 + lane;
    int peer_rr = 0;                // next peer to poll (warp 0)
    float th_g = 0.f;               // the group's published threshold, re-read once per chunk

#ifdef PK_PRE_CHUNK
    constexpr int kChunk = PK_PRE_CHUNK;
#else
    constexpr int kChunk = 32;
#endif
    static_assert(kChunk <= 32, "a chunk is tested one candidate per lane");
    // the claim of the NEXT chunk is issued before the current one is scored: the cursor of the dynamic schedule
    // lives in global memory (shared by the CTAs working on the group) and its latency must not stall the warp
    int c_next = 0;
    if (lane == 0) c_next = cursor_base + atomicAdd(cursor, kChunk);
    for (;;) {
        const int c0 = __shfl_sync(kFull, c_next, 0);
        if (c0 >= it.cand_end) break;
        if (lane == 0) c_next = cursor_base + atomicAdd(cursor, kChunk);
        const int c1 = min(c0 + kChunk, it.cand_end);
        uint64_t offv = 0;
        uint2 cm = make_uint2(0u, 0u);
        int myrow = c0 + lane;
        bool live = c0 + lane < c1;
        if (!STORE && live) {
            const int pos = c0 + lane;
            if (pos < P.n_list) myrow = __ldg(P.row_list + pos);
            else {
                myrow = P.cand_begin + (pos - P.n_list);
                if (P.skip != nullptr && __ldg(P.skip + myrow) != 0) live = false;      // already scanned as a heavy row
            }
        }
        if (live) {
            offv = __ldg(P.lrec_off + myrow);
            if (!STORE) cm = __ldg(P.lmask + myrow);
        }
        uint32_t surv = __ballot_sync(kFull, live);
        if (!STORE) {
            // thresholds published by the other instances of the group
            if (warp == 0 && P.n_peers > 0) {
                // one peer per chunk, round robin: its threshold and its rank-k_frac score for MY query, accepted only
                // under the same tag
                const size_t at = (size_t)pick[0] * 32 + lane;
                const unsigned long long pv = *reinterpret_cast<const volatile unsigned long long*>(P.peer_theta[peer_rr] + at);
                const unsigned long long ps = *reinterpret_cast<const volatile unsigned long long*>(P.peer_sfrac[peer_rr] + at);
                if ((unsigned int)(pv >> 32) == P.theta_tag) atomicMax(g_theta, pv);
                if ((unsigned int)(ps >> 32) == P.theta_tag) pcache[peer_rr * 32 + lane] = __uint_as_float((unsigned int)ps);
                peer_rr = (peer_rr + 1 == P.n_peers) ? 0 : peer_rr + 1;
                // the merged k-th best is at least the smallest rank-k_frac score over ALL ranks (0 = not known yet)
                const unsigned long long os = *reinterpret_cast<volatile unsigned long long*>(g_sfrac);
                float lo = ((unsigned int)(os >> 32) == P.theta_tag) ? __uint_as_float((unsigned int)os) : 0.f;
                for (int r = 0; r < P.n_peers; ++r) lo = fminf(lo, pcache[r * 32 + lane]);
                const float th2 = lo - P.theta_drop;
                if (th2 > 0.f) atomicMax(g_theta, ((unsigned long long)P.theta_tag << 32) | __float_as_uint(th2));
            }
            {
                const unsigned long long gv = *reinterpret_cast<volatile unsigned long long*>(g_theta);
                th_g = ((unsigned int)(gv >> 32) == P.theta_tag) ? __uint_as_float((unsigned int)gv) : 0.f;
            }
            if (th_g > reinterpret_cast<volatile float*>(theta)[lane]) {
                reinterpret_cast<volatile float*>(theta)[lane] = th_g;
                if (th_g > reinterpret_cast<volatile float*>(l1b)[lane])
                    l1_update(th_g, l1f + lane * 8, ub_text, inv_total, reinterpret_cast<volatile uint4*>(l1q) + lane,
                              reinterpret_cast<volatile float*>(l1b) + lane);
            }
            __syncwarp();
            // level 1, lane = candidate: can ANY query of the group still want it, judged by the number of jointly
            // filled text columns alone?  (8 bytes per candidate; two thirds of the candidates end here)
            bool any = false;
            const int nl = gd->n_lanes;
            #pragma unroll 4
            for (int q = 0; q < nl; ++q) {
                const uint4 qq = l1q[q];        // (re-read every chunk: the __syncwarp above orders it after the updates)
                const int j = min(__popc(qq.x & cm.x) + __popc(qq.y & cm.y), 63);
                const uint64_t adm = ((uint64_t)qq.w << 32) | qq.z;
                any = any || ((adm >> j) & 1ull);
            }
            surv &= __ballot_sync(kFull, any);
        }
        while (surv) {
            const int si = __ffs(surv) - 1;
            surv &= surv - 1;
            const int ci = STORE ? c0 + si : __shfl_sync(kFull, myrow, si);      // the candidate's store row
            const uint64_t off = __shfl_sync(kFull, offv, si);
            const uint4* cr = P.lrec + off;
            const uint32_t* cw = reinterpret_cast<const uint32_t*>(cr);
            const uint4 h0 = __ldg(cr), h1 = __ldg(cr + 1), h2 = __ldg(cr + 2), h3 = __ldg(cr + 3);
            // the first 32 club / friend ids sit at a fixed offset: their load does not wait for the header (words past
            // the sets, or past the record -- the buffer is padded -- are read and ignored)
            const uint32_t el0 = __ldg(cw + kLeanHdrWords + lane);
            const int c_uid = (int)h0.x, c_pub = (int)h0.y, c_gender = (int)h0.z, c_comp = (int)h0.w;
            const int c_age = (int)h1.x, c_r0 = (int)h1.y, c_r1 = (int)h1.z, c_r2 = (int)h1.w;
            const uint32_t c_ncl = h2.x, c_nfr = h2.y, c_ntok = h2.z;
            const int c_bcnt = (int)h2.w;
            const uint32_t o_clubs = kLeanHdrWords, o_friends = o_clubs + pad4(c_ncl), o_keys = o_friends + pad4(c_nfr),
                           o_w = o_keys + pad4(c_ntok);

            // the token keys / weights of a round are requested one round ahead (the first round here, before the
            // scalar fields and the sets), so that their latency is covered by other work
            const uint32_t* keys = cw + o_keys;
            const float* tokw = reinterpret_cast<const float*>(cw + o_w);
            uint32_t key_n = 0;
            float w_n = 0.f;
            if ((uint32_t)lane < c_ntok) { key_n = __ldg(keys + lane); w_n = __ldg(tokw + lane); }

            float sum = 0.f;
            int used = 0;
            // fixed scalar fields (src/recommender_similarity.cpp:38-77)
            if (q_pub >= 0 && c_pub >= 0) { const float4 f = N->fld[0]; sum += (q_pub == c_pub) ? f.w : f.z; ++used; }
            if (q_gender >= 0 && c_gender >= 0) { const float4 f = N->fld[1]; sum += (q_gender == c_gender) ? f.w : f.z; ++used; }
            if (q_comp > 0 && c_comp > 0) {
                float sg;
                if (q_comp < dim_comp && c_comp < dim_comp) sg = __ldg(P.tab_comp + c_comp * dim_comp + q_comp);
                else sg = lean_sigma(N->fld[2], (float)min(q_comp, c_comp) / (float)max(q_comp, c_comp));
                sum += sg; ++used;
            }
            if (q_age > 0 && c_age > 0) {
                float sg;
                if (q_age < dim_age && c_age < dim_age) sg = __ldg(P.tab_age + c_age * dim_age + q_age);
                else sg = lean_sigma(N->fld[3], (float)min(q_age, c_age) / (float)max(q_age, c_age));
                sum += sg; ++used;
            }
            if (q_acnt > 0 && c_bcnt > 0) {
                const int matches = (int)(q_r0 >= 0 && q_r0 == c_r0) + (int)(q_r1 >= 0 && q_r1 == c_r1) + (int)(q_r2 >= 0 && q_r2 == c_r2);
                sum += N->region[(q_acnt * 4 + c_bcnt) * 4 + matches];
                ++used;
            }
            // text columns present on both sides: every one contributes sigma(z(0)) unless a token matches
            const uint32_t b0 = q_m0 & h3.x, b1 = q_m1 & h3.y;
            const int joint = __popc(b0) + __popc(b1);

            // header-only upper bound against this query's threshold (PreParams): 2 S F / (S + F) >= theta with
            // S = s_ub / u, F = u / total  <=>  2 s_ub u / total >= theta (s_ub + u^2 / total)
            bool alive = q_active;
            if (!STORE) {
                const float th = fmaxf(reinterpret_cast<volatile float*>(theta)[lane], th_g);
                float s_ub = fmaf((float)joint, ub_text, sum);
                int u = used + joint;
                if (c_ncl > 0 && q_ncl > 0) { s_ub += ub_clubs; ++u; }
                if (c_nfr > 0 && q_nfr > 0) { s_ub += ub_friends; ++u; }
                const float fu = (float)u;
                alive = alive && (2.0f * s_ub * fu * inv_total >= th * fmaf(fu * fu, inv_total, s_ub));
                if (!__any_sync(kFull, alive)) continue;
            }

            sum += N->lut[0][b0 & 255u] + N->lut[1][(b0 >> 8) & 255u] + N->lut[2][(b0 >> 16) & 255u] + N->lut[3][b0 >> 24];
            if (q_m1 | h3.y)
                sum += N->lut[4][b1 & 255u] + N->lut[5][(b1 >> 8) & 255u] + N->lut[6][(b1 >> 16) & 255u] + N->lut[7][b1 >> 24];
            used += joint;

            // clubs / friends (:79-91): the record stores them back to back ([clubs padded to 4 | friends]) and the
            // group's table holds both sets ({id | set << 31, query mask}), so one round of <= 32 elements usually
            // covers both; a transpose of the hit masks gives every lane the hits of ITS query, counted by popc
            if (c_ncl + c_nfr > 0) {
                const uint32_t pcl = pad4(c_ncl), n_el = pcl + c_nfr;
                int inter_c = 0, inter_f = 0;
                for (uint32_t e0 = 0; e0 < n_el; e0 += 32) {
                    const uint32_t i = e0 + lane;
                    const bool is_fr = i >= pcl;
                    uint32_t m = 0;
                    if (is_fr ? (i < n_el) : (i < c_ncl)) {
                        const uint32_t key = ((e0 == 0u) ? el0 : __ldg(cw + o_clubs + i)) | (is_fr ? 0x80000000u : 0u);
                        const uint4 e1 = t_sets[hash_u32(key, log_cs)], e2 = t_sets[hash2_u32(key, log_cs)];
                        m = ((e1.x == key) ? e1.y : 0u) | ((e1.z == key) ? e1.w : 0u) | ((e2.x == key) ? e2.y : 0u) |
                            ((e2.z == key) ? e2.w : 0u);
                    }
                    const uint32_t hit_lanes = __ballot_sync(kFull, m != 0u);
                    if (hit_lanes == 0u) continue;
                    const uint32_t mine = transpose32(m, lane);
                    const uint32_t ncl_here = (pcl > e0) ? min(pcl - e0, 32u) : 0u;           // warp-uniform
                    const uint32_t cmask = (ncl_here >= 32u) ? 0xffffffffu : ((1u << ncl_here) - 1u);
                    inter_c += __popc(mine & cmask);
                    inter_f += __popc(mine & ~cmask);
                }
                if (c_ncl > 0 && q_ncl > 0) {
                    const float4 f = N->fld[5];
                    sum += (inter_c > 0) ? lean_sigma(f, (float)inter_c * q_rsc * __uint_as_float(h3.z)) : f.z;
                    ++used;
                }
                if (c_nfr > 0 && q_nfr > 0) {
                    const float4 f = N->fld[6];
                    sum += (inter_f > 0) ? lean_sigma(f, (float)inter_f * q_rsf * __uint_as_float(h3.w)) : f.z;
                    ++used;
                }
            }

            // text columns (:93-114): the lean weights are already divided by the cell norms
            if (c_ntok > 0) {
                float acc = 0.f;
                int cur = -1;                                         // per lane: column of the open cell of MY query
                for (uint32_t cb = 0; cb < c_ntok; cb += 32) {
                    const uint32_t i = cb + lane;
                    uint32_t m = 0, id = 0;
                    const uint32_t key = key_n;
                    const float w = w_n;
                    if (i + 32u < c_ntok) { key_n = __ldg(keys + i + 32u); w_n = __ldg(tokw + i + 32u); }
                    if (i < c_ntok) {
                        // 2-choice cuckoo probe: both buckets are read unconditionally (two independent LDS.128), a key
                        // lives in at most one slot and empty slots carry entry 0, so the matches can simply be OR-ed
                        const uint4 e1 = t_ck[hash_u32(key, log_cb)], e2 = t_ck[hash2_u32(key, log_cb)];
                        id = ((e1.x == key) ? e1.y : 0u) | ((e1.z == key) ? e1.w : 0u) | ((e2.x == key) ? e2.y : 0u) |
                             ((e2.z == key) ? e2.w : 0u);
                    }
                    const uint32_t hit_lanes = __ballot_sync(kFull, id != 0u);
                    if (hit_lanes == 0u) continue;
                    if (id) {
                        const uint2 en = t_ent[id - 1u];
                        m = en.x;
                        stage[lane] = make_uint4(m, en.y, __float_as_uint(w), key >> kTokBits);
                    }
                    // bit j of `mine` = "token j of this chunk belongs to MY query": each lane walks only its own hits
                    uint32_t mine = transpose32(m, lane);
                    if (!alive) mine = 0u;
                    __syncwarp();
                    while (mine) {
                        const int j = __ffs(mine) - 1;
                        mine &= mine - 1;
                        const uint4 h = stage[j];
                        if ((int)h.w != cur) {
                            if (acc != 0.f) { const float4 f = N->fld[kNumFixed + cur]; sum += lean_sigma(f, acc) - f.z; }
                            acc = 0.f;
                            cur = (int)h.w;
                        }
                        acc = fmaf(vals[h.y + __popc(h.x & lane_lt)], __uint_as_float(h.z), acc);
                    }
                    __syncwarp();
                }
                if (acc != 0.f) { const float4 f = N->fld[kNumFixed + cur]; sum += lean_sigma(f, acc) - f.z; }
            }

            float fas = 0.f;
            if (used > 0) {
                const float S = __fdividef(sum, (float)used);
                const float F = (float)used * inv_total;
                fas = __fdividef(2.0f * S * F, S + F);
                fas = fmaxf(fas, 0.f);
            }

            if (STORE) {
                P.out_f[(size_t)(it.out_slot + (ci - it.cand_begin)) * 32 + lane] = q_active ? fas : -1.0f;
            } else {
                const uint64_t key = make_key(__float_as_uint(fas), ci);
                const uint64_t tl = reinterpret_cast<volatile uint64_t*>(tau)[lane];
                // (a score below theta cannot be needed either: the list only has to hold what reaches a_k - 2 margin)
                uint32_t pass = __ballot_sync(kFull, alive && key > tl && fas >= reinterpret_cast<volatile float*>(theta)[lane]);
                if (pass) {
                    // exclusions only matter for the few candidates that reach a list
                    bool ok = true;
                    if (P.exclude_mode >= PK_EXCLUDE_SELF && c_uid == q_uid) ok = false;
                    if (P.exclude_mode >= PK_EXCLUDE_SELF_AND_ADJ) {
                        const uint32_t mx = probe_set(t_adj, log_sa, (uint32_t)c_uid);
                        if (mx & lanebit) ok = false;
                    }
                    pass &= __ballot_sync(kFull, ok);
                    while (pass) {
                        const int j = __ffs(pass) - 1;
                        pass &= pass - 1;
                        const uint64_t kj = __shfl_sync(kFull, key, j);
                        shortlist_insert(lists + j * Ks, Ks, kj, locks + j, reinterpret_cast<volatile uint64_t*>(tau) + j, lane, P.K,
                                         P.theta_drop, reinterpret_cast<volatile float*>(theta) + j, g_theta - lane + j, P.theta_tag,
                                         g_sfrac ? g_sfrac - lane + j : nullptr, P.k_frac,
                                         l1f + j * 8, ub_text, inv_total, reinterpret_cast<volatile uint4*>(l1q) + j,
                                         reinterpret_cast<volatile float*>(l1b) + j);
                    }
                }
            }
        }
    }

    if (!STORE) {
        __syncthreads();
        uint64_t* dst = P.out_keys + (size_t)it.out_slot * 32 * Ks;
        for (int i = tid; i < 32 * Ks; i += kPreThreads) dst[i] = lists[i];
        __syncthreads();                  // the lists and tables are re-used by the next round
    }
    }   // round
}

// Per query: is the merged shortlist (Ks approximate keys, descending) guaranteed to contain the exact
// top-k?  Yes when it holds every candidate (fewer than Ks entries) or when its last approximate score is
// more than 2*margin below the k-th one.  Also writes the shortlist's candidate rows for the exact pass.
struct CheckParams {
    const uint64_t* short_keys;   // [nq][Ks]
    const int32_t* q_group;       // group of each query, -1 = no profile
    const int32_t* q_lane;
    int32_t* cand_rows;           // [(group*32 + lane)*Ks + i], -1 = empty
    int32_t* q_flag;              // preset by the host (groups the pre-filter cannot take); OR-ed here
    int32_t* group_flag;
    const int32_t* lean_fail;     // [group] the group's pre-filter table was not built (group_prep_kernel): nothing is proven
    int nq, K, Ks;
    float margin;
    // the exact re-score as pairs of the lane-per-pair kernel: slot q * Ks + i = (query row, candidate row or -1, uid)
    int32_t *pa_row, *pb_row, *pkey_id;
    const int32_t* uid;
    const GroupDesc* groups;
    int own_lo, own_hi;           // query groups this rank re-scores
};

__global__ void shortlist_check_kernel(const CheckParams P) {
    const int q = blockIdx.x;
    const int g = P.q_group[q];
    if (g < 0) return;
    const uint64_t* keys = P.short_keys + (size_t)q * P.Ks;
    int32_t* rows = P.cand_rows + ((size_t)g * 32 + P.q_lane[q]) * P.Ks;
    const int qrow = P.groups[g].row[P.q_lane[q]];
    const bool own = g >= P.own_lo && g < P.own_hi;
    int32_t* pa = P.pa_row + (size_t)q * P.Ks;
    int32_t* pb = P.pb_row + (size_t)q * P.Ks;
    int32_t* pid = P.pkey_id + (size_t)q * P.Ks;
    if (keys[0] == kPoisonKey) {          // some rank could not build this group's fp32 tables (MergeParams): nothing is proven
        for (int i = threadIdx.x; i < P.Ks; i += blockDim.x) { rows[i] = -1; pa[i] = qrow; pb[i] = -1; pid[i] = 0; }
        if (threadIdx.x == 0) { P.q_flag[q] = 1; atomicOr(P.group_flag + g, 1); }
        return;
    }
    // only candidates whose approximate score can still reach the exact top-k (approx >= a_k - 2 margin) are re-scored
    const float a_k0 = __uint_as_float(key_score_bits(keys[P.K - 1]));     // 0 when the list holds fewer than k entries
    const float keep_from = a_k0 - 2.0f * P.margin - 1e-7f;
    for (int i = threadIdx.x; i < P.Ks; i += blockDim.x) {
        const uint64_t k = keys[i];
        const int r = (k && __uint_as_float(key_score_bits(k)) >= keep_from) ? key_id(k) : -1;
        rows[i] = r;
        pa[i] = qrow;
        pb[i] = own ? r : -1;
        pid[i] = r >= 0 ? P.uid[r] : 0;
    }
    if (threadIdx.x == 0) {
        const uint64_t last = keys[P.Ks - 1];
        bool complete = true;
        if (last != 0ull) {
            const float a_last = __uint_as_float(key_score_bits(last));
            const float a_k = __uint_as_float(key_score_bits(keys[P.K - 1]));
            complete = a_last < a_k - 2.0f * P.margin - 1e-7f;
        }
        if (P.lean_fail[g]) complete = false;
        if (!complete) { P.q_flag[q] = 1; atomicOr(P.group_flag + g, 1); }
    }
}

}  // namespace pk
