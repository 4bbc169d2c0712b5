// pk_clubs.cuh -- collaborative club propagation (reference src/recommender_clubs.cpp:10-73).
//
// The reference walks, per user, 1-hop entries (every friend f with w_f = FAS(user,f) > 0 adds w_f to
// each of f's clubs) and then 2-hop entries (every fof in adj[f] with s = FAS(f,fof) > 0 adds w_f*s to
// each of fof's clubs), skipping the user's own clubs, accumulating in double in visiting order.
// The FAS values come from the pair kernel (or the per-edge cache); this kernel does the scatter.  To keep every
// club's double sum in the reference's order (so the float cast is bit-identical) it is deterministic:
//   A  all threads, parallel over the edges: the distinct clubs of the user are hashed into a per-user table
//      (order-free) and every contribution gets {table slot, value} at its position in the visiting order;
//   B  every warp replays that flat stream and adds only into the table slots it owns (slot % warps): contributions to
//      different clubs commute, a club is always summed by one warp in stream order;
//   C  all threads: sort keys, compacted.
#pragma once
#include "pk_layout.h"

namespace pk {

#ifndef PK_CLUB_THREADS
#define PK_CLUB_THREADS 512
#endif
constexpr int kClubThreads = PK_CLUB_THREADS;

struct ClubEdge {
    int64_t w_idx;        // index of w_f in the scan output
    int64_t s_idx;        // index of FAS(f, fof), or -1 for a 1-hop entry
    int32_t target_row;   // whose clubs receive the contribution
    int32_t pad_;
};

struct ClubParams {
    const uint4* rec;
    const uint64_t* rec_off;
    const float* fas;
    const ClubEdge* edges;
    const uint32_t* con_off;          // per edge: first contribution index, relative to the user
    const int32_t* user_edge_begin;   // [nu+1]
    const int32_t* user_slot_begin;   // [nu+1] table slice of each user (power-of-two length)
    const int32_t* user_row;          // [nu]
    int32_t* con_slot;                // per contribution (indexed inside the user's slot slice): table slot or -1
    uint64_t* slot_key;               // in: 0; during: club id + 1; out: sort key of (score, club id) or 0
    double* slot_acc;                 // in: 0
    uint64_t* out_keys;               // the user's keys compacted to the front of its slice (selection reads only these)
    int32_t* out_cnt;                 // [nu] number of keys of each user
};

__global__ void __launch_bounds__(kClubThreads) clubs_scatter_kernel(const ClubParams P) {
    const int u = blockIdx.x, tid = threadIdx.x;
    const int eb = P.user_edge_begin[u], ee = P.user_edge_begin[u + 1];
    if (eb == ee) { if (tid == 0) P.out_cnt[u] = 0; return; }     // no entries (unknown user, no profiled friend)
    const int sb = P.user_slot_begin[u];
    const uint32_t cap = (uint32_t)(P.user_slot_begin[u + 1] - sb);
    const uint32_t msk = cap - 1u;
    int log_cap = 0;
    while ((1u << log_cap) < cap) ++log_cap;
    const uint32_t* ur = reinterpret_cast<const uint32_t*>(P.rec + P.rec_off[P.user_row[u]]);
    const uint32_t n_ucl = ur[8];
    const uint32_t* ucl = ur + kHdrWords;
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(P.slot_key + sb);
    double* acc = P.slot_acc + sb;
    int32_t* cslot = P.con_slot + sb;
    // value of each contribution: lives in the user's slice of out_keys until phase C overwrites it with the keys
    // (a user has at most cap / 2 contributions)
    double* cval = reinterpret_cast<double*>(P.out_keys + sb);
    uint32_t n_con;
    {
        const ClubEdge el = P.edges[ee - 1];
        n_con = P.con_off[ee - 1] + reinterpret_cast<const uint32_t*>(P.rec + P.rec_off[el.target_row])[8];
    }

    // phase A (parallel over edges): validity of each entry, distinct clubs into the table
    for (int e = eb + tid; e < ee; e += kClubThreads) {
        const ClubEdge ed = P.edges[e];
        const double w = (double)P.fas[ed.w_idx];
        bool valid = w > 0.0;                                     // src/recommender_clubs.cpp:38,51
        if (ed.s_idx >= 0) valid = valid && ((double)P.fas[ed.s_idx] > 0.0);   // :57
        const uint32_t* tr = reinterpret_cast<const uint32_t*>(P.rec + P.rec_off[ed.target_row]);
        const uint32_t ncl = tr[8];
        const uint32_t c0 = P.con_off[e];
        double v = w;
        if (ed.s_idx >= 0) v = __dmul_rn(w, (double)P.fas[ed.s_idx]);       // contrib = wuf * s_f_fof (:58)
        for (uint32_t j = 0; j < ncl; ++j) {
            const uint32_t club = tr[kHdrWords + j];
            bool skip = !valid;
            for (uint32_t i = 0; i < n_ucl && !skip; ++i) skip = (ucl[i] == club);   // :40,60 (user's own clubs)
            int32_t sl = -1;
            if (!skip) {
                uint32_t slot = (cap > 1u) ? ((club * 0x9E3779B1u) >> (32 - log_cap)) : 0u;
                const unsigned long long want = (unsigned long long)club + 1ull;
                for (;;) {
                    unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(keys + slot);
                    if (cur == 0ull) {
                        const unsigned long long old = atomicCAS(keys + slot, 0ull, want);
                        if (old == 0ull) { sl = (int32_t)slot; break; }
                        cur = old;
                    }
                    if (cur == want) { sl = (int32_t)slot; break; }
                    slot = (slot + 1) & msk;
                }
            }
            cslot[c0 + j] = sl;
            cval[c0 + j] = v;
        }
    }
    __syncthreads();

    // phase B: the ordered stream is replayed by ALL warps, warp w adding only into the table slots it owns
    // (slot % warps == w): every club is still summed by one warp in stream order, different clubs proceed in parallel.
    // A warp scans the whole stream 32 contributions at a time (the next chunk's loads in flight), COMPACTS its own
    // contributions into a 64-entry ring in shared memory and adds them 32 at a time; contributions to the same slot
    // inside such a batch are added by the lowest lane, in order.  (Round 1 let every THREAD replay the whole stream;
    // the first rewrite of round 2 used one warp edge by edge -- ~370 dependent L2 round trips per user, 9 ms per
    // 4096-user batch; one warp over the flat stream took 3.3 ms, bound by the users with the most entries: 8000
    // edges = 1200 serial read-modify-write steps.)
    constexpr int kWarps = kClubThreads / 32;
    __shared__ int32_t ring_sl[kWarps][64];
    __shared__ double ring_v[kWarps][64];
    {
        const int lane = tid & 31, warp = tid >> 5;
        const uint32_t lane_lt = (1u << lane) - 1u;
        int32_t* rsl = ring_sl[warp];
        double* rv = ring_v[warp];
        int cnt = 0, done = 0;                                    // warp-uniform
        auto add_batch = [&](int n) {
            int32_t sl = -1;
            if (lane < n) sl = rsl[(done + lane) & 63];
            const uint32_t live = __ballot_sync(0xffffffffu, sl >= 0);
            const uint32_t same = __match_any_sync(0xffffffffu, sl >= 0 ? sl : -1 - lane) & live;
            if (sl >= 0 && (__ffs(same) - 1) == lane) {
                double a = acc[sl];
                for (uint32_t mm = same; mm; mm &= mm - 1u) a = __dadd_rn(a, rv[(done + __ffs(mm) - 1) & 63]);
                acc[sl] = a;
            }
            __syncwarp();
        };
        int32_t sl_n = -1;
        double v_n = 0.0;
        if ((uint32_t)lane < n_con) { sl_n = cslot[lane]; v_n = cval[lane]; }
        for (uint32_t c0 = 0; c0 < n_con; c0 += 32) {
            const int32_t sl = sl_n;
            const double v = v_n;
            const uint32_t cn = c0 + 32u + lane;
            sl_n = -1;
            if (cn < n_con) { sl_n = cslot[cn]; v_n = cval[cn]; }
            const bool mine = sl >= 0 && (sl & (kWarps - 1)) == warp;
            const uint32_t mm = __ballot_sync(0xffffffffu, mine);
            if (mm == 0u) continue;
            if (mine) {
                const int pos = (cnt + __popc(mm & lane_lt)) & 63;
                rsl[pos] = sl;
                rv[pos] = v;
            }
            cnt += __popc(mm);
            __syncwarp();
            if (cnt - done >= 32) { add_batch(32); done += 32; }
        }
        if (cnt > done) add_batch(cnt - done);
    }
    __syncthreads();

    // phase C: sort keys (score desc, club id asc; :67-70), compacted to the front of the user's slice
    __shared__ int n_keys;
    if (tid == 0) n_keys = 0;
    __syncthreads();
    for (uint32_t i0 = 0; i0 < cap; i0 += kClubThreads) {
        const uint32_t i = i0 + tid;
        const unsigned long long c1 = (i < cap) ? keys[i] : 0ull;
        const bool has = c1 != 0ull;
        const uint32_t m = __ballot_sync(0xffffffffu, has);
        int base = 0;
        if ((tid & 31) == 0 && m) base = atomicAdd(&n_keys, __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (has) P.out_keys[sb + base + __popc(m & ((1u << (tid & 31)) - 1u))] =
            make_key(__float_as_uint((float)acc[i]), (int32_t)(uint32_t)(c1 - 1ull));
    }
    __syncthreads();
    if (tid == 0) P.out_cnt[u] = n_keys;
}

}  // namespace pk
