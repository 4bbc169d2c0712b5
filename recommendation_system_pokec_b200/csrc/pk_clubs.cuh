// pk_clubs.cuh -- collaborative club propagation (reference src/recommender_clubs.cpp:10-73).
//
// The reference walks, per user, 1-hop entries (every friend f with w_f = FAS(user,f) > 0 adds w_f to
// each of f's clubs) and then 2-hop entries (every fof in adj[f] with s = FAS(f,fof) > 0 adds w_f*s to
// each of fof's clubs), skipping the user's own clubs, accumulating in double in visiting order.
// The FAS values come from the store-mode scan; this kernel does the scatter.  To keep every club's
// double sum in the reference's order (so the float cast is bit-identical), the scatter is made
// deterministic: the distinct clubs of a user are first hashed into a per-user table (order-free),
// then thread t of the CTA replays the ordered contribution stream and adds only into the slots it
// owns (slot % blockDim == t).  Different clubs proceed in parallel, one club is always summed by one
// thread in stream order.
#pragma once
#include "pk_layout.h"

namespace pk {

constexpr int kClubThreads = 256;

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
};

__global__ void __launch_bounds__(kClubThreads) clubs_scatter_kernel(const ClubParams P) {
    const int u = blockIdx.x, tid = threadIdx.x;
    const int eb = P.user_edge_begin[u], ee = P.user_edge_begin[u + 1];
    if (eb == ee) return;                 // no entries (unknown user, no profiled friend): the slice stays empty
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

    // phase A (parallel over edges): validity of each entry, distinct clubs into the table
    for (int e = eb + tid; e < ee; e += kClubThreads) {
        const ClubEdge ed = P.edges[e];
        const double w = (double)P.fas[ed.w_idx];
        bool valid = w > 0.0;                                     // src/recommender_clubs.cpp:38,51
        if (ed.s_idx >= 0) valid = valid && ((double)P.fas[ed.s_idx] > 0.0);   // :57
        const uint32_t* tr = reinterpret_cast<const uint32_t*>(P.rec + P.rec_off[ed.target_row]);
        const uint32_t ncl = tr[8];
        const uint32_t c0 = P.con_off[e];
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
        }
    }
    __syncthreads();

    // phase B (every thread replays the ordered stream, adds into the slots it owns)
    for (int e = eb; e < ee; ++e) {
        const ClubEdge ed = P.edges[e];
        const double w = (double)P.fas[ed.w_idx];
        double v = w;
        if (ed.s_idx >= 0) v = __dmul_rn(w, (double)P.fas[ed.s_idx]);   // contrib = wuf * s_f_fof (:58)
        const uint32_t c0 = P.con_off[e];
        const uint32_t c1 = (e + 1 < ee) ? P.con_off[e + 1] : 0xffffffffu;
        const uint32_t* tr = reinterpret_cast<const uint32_t*>(P.rec + P.rec_off[ed.target_row]);
        const uint32_t ncl = (c1 != 0xffffffffu) ? (c1 - c0) : tr[8];
        for (uint32_t j = 0; j < ncl; ++j) {
            const int32_t sl = cslot[c0 + j];
            if (sl >= 0 && (sl & (kClubThreads - 1)) == tid) acc[sl] = __dadd_rn(acc[sl], v);
        }
    }
    __syncthreads();

    // phase C: sort keys (score desc, club id asc; :67-70)
    for (uint32_t i = tid; i < cap; i += kClubThreads) {
        const unsigned long long c1 = keys[i];
        if (c1 != 0ull) keys[i] = make_key(__float_as_uint((float)acc[i]), (int32_t)(uint32_t)(c1 - 1ull));
    }
}

}  // namespace pk
