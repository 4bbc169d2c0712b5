// pk_pairs.cuh -- exact FAS of explicit (A, B) pairs, one LANE per pair.
//
// fas_scan_kernel (pk_scan.cuh) shares every streamed candidate between the <= 32 queries of a group through
// inverted tables in shared memory.  That pays when many queries meet the same candidates (full scan, the
// friends x candidates matrix of recommend_collaborative); for lists of unrelated pairs it wastes the warp --
// the per-edge FAS cache scored a 32 x 32 block per 32 wanted pairs, recommend_graph_registration ran 1-lane
// groups.  Here a thread owns one pair and walks the two packed records itself: set overlaps by merging the
// (sorted) id lists, tf-idf dot products by merging the token keys of the jointly filled columns.  Same
// arithmetic, same rounding points and the same order of additions as fas_scan_kernel, i.e. as
// Recommender::profile_similarity (reference src/recommender_similarity.cpp:10-124, src/recommender.cpp:68-139).
#pragma once
#include "pk_scan.cuh"

namespace pk {

struct PairParams {
    const uint4* rec;
    const uint64_t* rec_off;
    const NormTables* norm;
    const double* tab_comp;      // [dim][dim] sigmoid(z(min/max)), index [B][A]
    const double* tab_age;
    const int32_t* a_row;        // per pair: store row of A (the query side of profile_similarity(A, B)); < 0 = no profile
    const int32_t* b_row;
    int64_t n;
    int T, nf;                   // nf = 3 + T + 2 (FIELDS layout of fas_scan_kernel)
    float* out_f;                // [n] FAS, -1 where a side has no profile
    float* out_fields;           // FIELDS: [n][nf]
    // optional epilogue: sort keys (score, id) for the list modes -- key_id[i] is the id the key carries
    const int32_t* key_id;
    uint64_t* out_keys;
};

// |{b in B : b in set(A)}| for two ascending id lists (duplicates allowed on either side; src/recommender.cpp:119-128
// builds a set of A and walks B with multiplicity).  One flat two-pointer loop (lanes of a warp stay in the same
// loop whatever their lengths); binary search when one side is much longer.
__device__ __forceinline__ int overlap_sorted(const uint32_t* __restrict__ A, uint32_t nA, const uint32_t* __restrict__ B,
                                              uint32_t nB) {
    int cnt = 0;
    if (nA >= 8u * nB) {                       // few b's against a long A: binary search each b
        for (uint32_t j = 0; j < nB; ++j) {
            const uint32_t b = __ldg(B + j);
            uint32_t lo = 0, hi = nA;
            while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (__ldg(A + mid) < b) lo = mid + 1; else hi = mid; }
            cnt += (lo < nA && __ldg(A + lo) == b) ? 1 : 0;
        }
    } else if (nB >= 8u * nA) {                // few distinct a's against a long B: count each a's run in B
        uint32_t prev = 0;
        for (uint32_t i = 0; i < nA; ++i) {
            const uint32_t a = __ldg(A + i);
            if (i > 0 && a == prev) continue;
            prev = a;
            uint32_t lo = 0, hi = nB;
            while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (__ldg(B + mid) < a) lo = mid + 1; else hi = mid; }
            uint32_t lo2 = lo, hi2 = nB;
            while (lo2 < hi2) { const uint32_t mid = (lo2 + hi2) >> 1; if (__ldg(B + mid) <= a) lo2 = mid + 1; else hi2 = mid; }
            cnt += (int)(lo2 - lo);
        }
    } else {
        // (measured on the collaborative batch of the bench, 72.9 M pairs: plain loop 30.1 ms, one element of look-ahead
        //  on both sides 28.9 ms, four elements per 16-byte load selected from a register window 58.6 ms)
        uint32_t i = 0, j = 0;
#ifndef PK_PAIR_NO_LOOKAHEAD
        uint32_t a = __ldg(A), b = __ldg(B), a1 = __ldg(A + 1), b1 = __ldg(B + 1);
        while (i < nA && j < nB) {
            cnt += (a == b) ? 1 : 0;
            if (a < b) { ++i; a = a1; a1 = __ldg(A + i + 1); }
            else { ++j; b = b1; b1 = __ldg(B + j + 1); }
        }
#else
        while (i < nA && j < nB) {
            const uint32_t a = __ldg(A + i), b = __ldg(B + j);
            cnt += (a == b) ? 1 : 0;           // this b is a member; the next b may equal a again, so only j moves
            i += (a < b) ? 1u : 0u;
            j += (a >= b) ? 1u : 0u;
        }
#endif
    }
    return cnt;
}

// The sigma(z(s)) chain -- an fp64 division, exp and another division -- is what a pair spends most instructions
// on, and only a few fields per pair need it (those with a non-zero similarity).  Executed where it arises it runs
// with a handful of lanes active.  So a lane QUEUES its pending fields {field, numerator, denominator} in order
// (clubs, friends, matched text columns ascending), the warp evaluates queue entry k of every lane together, and
// the lane then adds sigma values in the reference's field order, taking sigma(z(0)) for the joint fields that are
// not in the queue.
constexpr int kPairQueue = 2 + kMaxT;

template <bool FIELDS>
__global__ void __launch_bounds__(128) fas_pairs_kernel(const PairParams P) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P.n) return;
    const NormTables* N = P.norm;
    const int ra = P.a_row[p], rb = P.b_row[p];
    float* fo = nullptr;
    if (FIELDS) {
        fo = P.out_fields + (size_t)p * P.nf;
        for (int j = 0; j < P.nf; ++j) fo[j] = -1.0f;
    }
    if (ra < 0 || rb < 0) {
        P.out_f[p] = -1.0f;
        if (P.out_keys) P.out_keys[p] = 0ull;
        return;
    }
    const uint4* ar = P.rec + P.rec_off[ra];
    const uint4* br = P.rec + P.rec_off[rb];
    const uint32_t* aw = reinterpret_cast<const uint32_t*>(ar);
    const uint32_t* bw = reinterpret_cast<const uint32_t*>(br);
    const uint4 a0 = __ldg(ar), a1 = __ldg(ar + 1), a2 = __ldg(ar + 2), a3 = __ldg(ar + 3), a4 = __ldg(ar + 4);
    const uint4 b0 = __ldg(br), b1 = __ldg(br + 1), b2 = __ldg(br + 2), b3 = __ldg(br + 3), b4 = __ldg(br + 4);
    const int q_pub = (int)a0.y, q_gender = (int)a0.z, q_comp = (int)a0.w, q_age = (int)a1.x;
    const int c_pub = (int)b0.y, c_gender = (int)b0.z, c_comp = (int)b0.w, c_age = (int)b1.x;
    const uint32_t q_ncl = a2.x, q_nfr = a2.y, q_ncells = a2.z, q_ntok = a2.w;
    const uint32_t c_ncl = b2.x, c_nfr = b2.y, c_ncells = b2.z, c_ntok = b2.w;
    const uint64_t q_mask = (uint64_t)a3.x | ((uint64_t)a3.y << 32), c_mask = (uint64_t)b3.x | ((uint64_t)b3.y << 32);
    const int q_acnt = (int)a3.z, c_bcnt = (int)b3.z;
    const int dim_comp = N->dim_comp, dim_age = N->dim_age;

    double sum = 0.0;
    int used = 0;
    // public / gender (src/recommender_similarity.cpp:38-50)
    if (q_pub >= 0 && c_pub >= 0) { sum += (q_pub == c_pub) ? N->sig1[0] : N->sig0[0]; ++used; }
    if (q_gender >= 0 && c_gender >= 0) { sum += (q_gender == c_gender) ? N->sig1[1] : N->sig0[1]; ++used; }
    // completion / age (:52-68)
    if (q_comp > 0 && c_comp > 0) {
        double sg;
        if (q_comp < dim_comp && c_comp < dim_comp) sg = __ldg(P.tab_comp + (size_t)c_comp * dim_comp + q_comp);
        else sg = pk_sigmoid(pk_z(N, 2, (double)min(q_comp, c_comp) / (double)max(q_comp, c_comp)));
        sum += sg; ++used;
    }
    if (q_age > 0 && c_age > 0) {
        double sg;
        if (q_age < dim_age && c_age < dim_age) sg = __ldg(P.tab_age + (size_t)c_age * dim_age + q_age);
        else sg = pk_sigmoid(pk_z(N, 3, (double)min(q_age, c_age) / (double)max(q_age, c_age)));
        sum += sg; ++used;
    }
    // region (:70-77, src/recommender.cpp:130-139)
    if (q_acnt > 0 && c_bcnt > 0) {
        const int q_r0 = (int)a1.y, q_r1 = (int)a1.z, q_r2 = (int)a1.w, c_r0 = (int)b1.y, c_r1 = (int)b1.z, c_r2 = (int)b1.w;
        const int matches = (int)(q_r0 >= 0 && c_r0 >= 0 && q_r0 == c_r0) + (int)(q_r1 >= 0 && c_r1 >= 0 && q_r1 == c_r1) +
                            (int)(q_r2 >= 0 && c_r2 >= 0 && q_r2 == c_r2);
        sum += N->region[(q_acnt * 4 + c_bcnt) * 4 + matches];
        ++used;
        if (FIELDS) fo[0] = (float)((double)matches / (sqrt((double)q_acnt) * sqrt((double)c_bcnt)));
    }
    const uint32_t qo_clubs = kHdrWords, qo_friends = qo_clubs + pad4(q_ncl), qo_cells = qo_friends + pad4(q_nfr),
                   qo_keys = qo_cells + 4 * q_ncells, qo_w = qo_keys + pad4(q_ntok);
    const uint32_t co_clubs = kHdrWords, co_friends = co_clubs + pad4(c_ncl), co_cells = co_friends + pad4(c_nfr),
                   co_keys = co_cells + 4 * c_ncells, co_w = co_keys + pad4(c_ntok);

    // ---- the pending fields of this pair: q_f = field index, q_num / q_den = similarity numerator / denominator;
    //      after the evaluation pass q_num holds sigma(z(s)) ----
    double q_num[kPairQueue], q_den[2];       // (only the two set fields carry a denominator; a text column's comes from its cells)
    uint8_t q_f[kPairQueue];
    int qn = 0;

    // clubs / friends (:79-91, src/recommender.cpp:119-128); the packed lists are ascending
    const bool has_c = q_ncl > 0 && c_ncl > 0, has_f = q_nfr > 0 && c_nfr > 0;
    int inter_c = 0, inter_f = 0;
    if (has_c) inter_c = overlap_sorted(aw + qo_clubs, q_ncl, bw + co_clubs, c_ncl);
    if (has_f) inter_f = overlap_sorted(aw + qo_friends, q_nfr, bw + co_friends, c_nfr);
    if (inter_c > 0) {
        q_f[qn] = 5; q_num[qn] = (double)inter_c;
        q_den[qn] = __hiloint2double((int)a4.y, (int)a4.x) * __hiloint2double((int)b4.y, (int)b4.x); ++qn;
    }
    if (inter_f > 0) {
        q_f[qn] = 6; q_num[qn] = (double)inter_f;
        q_den[qn] = __hiloint2double((int)a4.w, (int)a4.z) * __hiloint2double((int)b4.w, (int)b4.z); ++qn;
    }
    if (FIELDS) {
        if (has_c) { fo[1] = 0.0f; fo[3 + P.T] = __int_as_float(inter_c); }
        if (has_f) { fo[2] = 0.0f; fo[3 + P.T + 1] = __int_as_float(inter_f); }
    }

    // text columns (:93-114): the token keys (column << 26 | id) of a record ascend over the whole record, so ONE
    // merge of the two key arrays meets every common token of every column, in ascending token order inside a column
    // (the order fas_scan_kernel adds the products in).  A column's dot product is queued when the next column starts.
    const uint64_t joint = q_mask & c_mask;
    used += (has_c ? 1 : 0) + (has_f ? 1 : 0) + __popcll(joint);
    uint64_t matched = 0ull;
    if (joint != 0ull) {
        const double* qwt = reinterpret_cast<const double*>(aw + qo_w);
        const double* cwt = reinterpret_cast<const double*>(bw + co_w);
        const uint32_t* ka_p = aw + qo_keys;
        const uint32_t* kb_p = bw + co_keys;
        uint32_t i = 0, j = 0;
        int cur = -1;
        double dot = 0.0;
#ifndef PK_PAIR_NO_LOOKAHEAD
        uint32_t ka = __ldg(ka_p), kb = __ldg(kb_p), ka1 = __ldg(ka_p + 1), kb1 = __ldg(kb_p + 1);
#endif
        while (i < q_ntok && j < c_ntok) {
#ifdef PK_PAIR_NO_LOOKAHEAD
            const uint32_t ka = __ldg(ka_p + i), kb = __ldg(kb_p + j);
#endif
            if (ka == kb) {
                const int ct = (int)(ka >> kTokBits);
                if (ct != cur) {
                    if (dot != 0.0) { q_f[qn] = (uint8_t)(kNumFixed + cur); q_num[qn] = dot; ++qn; matched |= 1ull << cur; }
                    cur = ct; dot = 0.0;
                }
                dot = __dadd_rn(dot, __dmul_rn(__ldg(qwt + i), __ldg(cwt + j)));
            }
            const bool adv_a = ka <= kb, adv_b = kb <= ka;
#ifndef PK_PAIR_NO_LOOKAHEAD
            if (adv_a) { ++i; ka = ka1; ka1 = __ldg(ka_p + i + 1); }
            if (adv_b) { ++j; kb = kb1; kb1 = __ldg(kb_p + j + 1); }
#else
            i += adv_a ? 1u : 0u;
            j += adv_b ? 1u : 0u;
#endif
        }
        if (dot != 0.0) { q_f[qn] = (uint8_t)(kNumFixed + cur); q_num[qn] = dot; ++qn; matched |= 1ull << cur; }
    }

    // ---- evaluation pass: entry k of every lane of the warp together ----
    for (int k = 0; k < qn; ++k) {
        const int f = q_f[k];
        double den = 0.0;
        if (f >= kNumFixed) {                    // text: the two cell norms (cells are indexed by the rank of the column)
            const int ct = f - kNumFixed;
            const uint64_t below = (1ull << ct) - 1ull;
            const uint4 qc = __ldg(ar + (qo_cells >> 2) + __popcll(q_mask & below));
            const uint4 cc = __ldg(br + (co_cells >> 2) + __popcll(c_mask & below));
            den = __hiloint2double((int)qc.w, (int)qc.z) * __hiloint2double((int)cc.w, (int)cc.z);
        } else den = q_den[k & 1];
        const float s = (den > 0.0) ? (float)(q_num[k] / den) : 0.0f;
        q_num[k] = (f >= kNumFixed && s == 1.0f) ? N->sig1[f] : pk_sigmoid(pk_z(N, f, (double)s));
        if (FIELDS) fo[f >= kNumFixed ? 3 + (f - kNumFixed) : f - 4] = s;
    }

    // ---- sum in the reference's field order: clubs, friends, text columns ascending ----
    int k = 0;
    if (has_c) sum += (inter_c > 0) ? q_num[k++] : N->sig0[5];
    if (has_f) sum += (inter_f > 0) ? q_num[k++] : N->sig0[6];
    uint64_t jm = joint;
    while (jm) {
        const int ct = __ffsll((long long)jm) - 1;
        jm &= jm - 1;
        if ((matched >> ct) & 1ull) sum += q_num[k++];
        else {
            sum += N->sig0[kNumFixed + ct];
            if (FIELDS) fo[3 + ct] = 0.0f;
        }
    }
    // S, F, harmonic mean (:116-123)
    float fas = 0.0f;
    if (used > 0) {
        const double S = sum / (double)used;
        const double F = N->ftab[used];
        fas = (float)((2.0 * S * F) / (S + F));
    }
    P.out_f[p] = fas;
    if (P.out_keys) P.out_keys[p] = make_key(__float_as_uint(fas), P.key_id[p]);
}

}  // namespace pk
