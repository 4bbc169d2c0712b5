// fas_oracle.cpp -- TEST INFRASTRUCTURE: CPU restatement of the reference's scoring hot path.
//
// This file restates, on flat sorted arrays instead of the reference's hash maps, the algorithm of
//   src/recommender_similarity.cpp:10-128   (FAS = profile_similarity)
//   src/recommender.cpp:43-163              (idf, tf-idf cosine, set / region similarity, raw cosine)
//   src/recommender_graph.cpp:10-227        (candidate gather, graph/interest, collaborative)
//   src/recommender_clubs.cpp:10-73         (collaborative clubs)
//   src/tfidf_index.cpp:7-30                (double-log idf variant)
// of pymlex/recommendation-system-pokec.  Each function cites the lines it follows.
//
// It is NOT part of the product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load it.  It is pinned against the compiled reference itself
// (oracle/_ref/libpokec_ref.so, built from /root/reference by oracle/build_oracle.py) in
// tests/test_oracle_cpu.py and against the committed fixtures in tests/golden/.
//
// Known, documented deviation: double sums over hash-map iteration order (src/recommender.cpp:75-113)
// are taken in ascending token order here.  With >= 3 addends the double can differ in the last bit,
// which is visible after the (float) cast with probability ~1e-8 per value.
//
// Build: g++ -O2 -ffp-contract=off -shared -fPIC (no FMA contraction: the reference is built for
// baseline x86-64, i.e. separately rounded multiply and add).
#include "pokec_world.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <unordered_map>
#include <utility>
#include <vector>

namespace {

struct IdfCol {
    bool present = false;            // idf_per_col has an entry for this column name
    std::vector<int32_t> token;      // sorted
    std::vector<float> idf;
    std::vector<int32_t> df;
    const float* find(int32_t tok) const {
        auto it = std::lower_bound(token.begin(), token.end(), tok);
        if (it == token.end() || *it != tok) return nullptr;
        return &idf[it - token.begin()];
    }
};

struct Oracle {
    pk_world w;  // pointers into copies below
    int n = 0, T = 0;
    std::vector<int32_t> uid, pub, comp, gender, age, region, tok_id, tok_tf, adj_uid, adj;
    std::vector<uint32_t> clubs, friends;
    std::vector<int64_t> clubs_ptr, friends_ptr, tok_ptr, adj_ptr;
    std::vector<float> norm_mean, norm_sd;
    std::vector<uint8_t> norm_has;
    std::unordered_map<int32_t, int32_t> row_of;      // uid -> profile row
    std::unordered_map<int32_t, int32_t> adj_row_of;  // uid -> adjacency row
    std::vector<IdfCol> idf;
    // sorted copies used for membership tests (the reference uses hash sets)
    std::vector<uint32_t> clubs_sorted, friends_sorted;

    int row(int32_t u) const {
        auto it = row_of.find(u);
        return it == row_of.end() ? -1 : it->second;
    }
    int adj_row(int32_t u) const {
        auto it = adj_row_of.find(u);
        return it == adj_row_of.end() ? -1 : it->second;
    }
};

template <class F>
void parallel_for(int64_t n, int nthreads, F f) {
    if (nthreads <= 1 || n <= 1) {
        for (int64_t i = 0; i < n; ++i) f(i);
        return;
    }
    std::atomic<int64_t> next(0);
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t)
        th.emplace_back([&] {
            for (;;) {
                int64_t i = next.fetch_add(1);
                if (i >= n) break;
                f(i);
            }
        });
    for (auto& t : th) t.join();
}

// ---- per-field similarities ---------------------------------------------------------------------

// src/recommender.cpp:119-128 vec_set_similarity: |{b in B : b in keys(A)}| / (sqrt|A| sqrt|B|),
// B counted with multiplicity, raw sizes in the denominator, result rounded to float.
// `As` is A sorted (duplicates harmless for membership).
float set_similarity(const uint32_t* As, int64_t na, const uint32_t* B, int64_t nb, int* inter_out) {
    if (inter_out) *inter_out = 0;
    if (na == 0 || nb == 0) return 0.0f;
    int inter = 0;
    for (int64_t j = 0; j < nb; ++j)
        if (std::binary_search(As, As + na, B[j])) ++inter;
    if (inter_out) *inter_out = inter;
    double denom = std::sqrt((double)na) * std::sqrt((double)nb);
    if (denom <= 0.0) return 0.0f;
    return (float)((double)inter / denom);
}

// src/recommender.cpp:130-139 region_similarity_local
float region_similarity(const int32_t* A, const int32_t* B) {
    int a_cnt = 0, b_cnt = 0, matches = 0;
    for (int i = 0; i < 3; ++i) {
        if (A[i] >= 0) ++a_cnt;
        if (B[i] >= 0) ++b_cnt;
        if (A[i] >= 0 && B[i] >= 0 && A[i] == B[i]) ++matches;
    }
    if (a_cnt == 0 || b_cnt == 0) return 0.0f;
    return (float)((double)matches / (std::sqrt((double)a_cnt) * std::sqrt((double)b_cnt)));
}

// src/recommender.cpp:68-117 tfidf_cosine_for_column.  Cells are sorted by token id.
// idf of a token missing from the column's table is 1.0f (:78).
float tfidf_cosine(const int32_t* ta, const int32_t* fa, int64_t na_, const int32_t* tb, const int32_t* fb,
                   int64_t nb_, const IdfCol& col) {
    if (na_ == 0 || nb_ == 0) return 0.0f;
    double dot = 0.0, na = 0.0, nb = 0.0;
    for (int64_t i = 0; i < na_; ++i) {
        const float* p = col.find(ta[i]);
        double idf = p ? *p : 1.0f;
        double wA = (double)fa[i] * idf;
        na += wA * wA;
    }
    for (int64_t j = 0; j < nb_; ++j) {
        const float* p = col.find(tb[j]);
        double idf = p ? *p : 1.0f;
        double wB = (double)fb[j] * idf;
        nb += wB * wB;
    }
    int64_t i = 0, j = 0;
    while (i < na_ && j < nb_) {
        if (ta[i] < tb[j]) ++i;
        else if (ta[i] > tb[j]) ++j;
        else {
            const float* p = col.find(ta[i]);
            double idf = p ? *p : 1.0f;
            double wA = (double)fa[i] * idf;
            double wB = (double)fb[j] * idf;
            dot += wA * wB;
            ++i; ++j;
        }
    }
    double denom = std::sqrt(na) * std::sqrt(nb);
    if (denom <= 0.0) return 0.0f;
    return (float)(dot / denom);
}

// src/recommender.cpp:141-163 cosine_counts_maps_local (used when the column has no idf entry)
float raw_cosine(const int32_t* ta, const int32_t* fa, int64_t na_, const int32_t* tb, const int32_t* fb,
                 int64_t nb_) {
    if (na_ == 0 || nb_ == 0) return 0.0f;
    double dot = 0.0, suma2 = 0.0, sumb2 = 0.0;
    for (int64_t i = 0; i < na_; ++i) suma2 += (double)fa[i] * fa[i];
    for (int64_t j = 0; j < nb_; ++j) sumb2 += (double)fb[j] * fb[j];
    if (suma2 <= 0.0 || sumb2 <= 0.0) return 0.0f;
    int64_t i = 0, j = 0;
    while (i < na_ && j < nb_) {
        if (ta[i] < tb[j]) ++i;
        else if (ta[i] > tb[j]) ++j;
        else { dot += (double)fa[i] * fb[j]; ++i; ++j; }
    }
    double norm = std::sqrt(suma2) * std::sqrt(sumb2);
    if (norm <= 0.0) return 0.0f;
    return (float)(dot / norm);
}

// src/recommender_similarity.cpp:18-26
inline double sigmoid(double x) {
    if (x >= 0) { double e = std::exp(-x); return 1.0 / (1.0 + e); }
    double e = std::exp(x);
    return e / (1.0 + e);
}

// src/recommender_similarity.cpp:28-36 / :105-111
inline double z_of(const Oracle& o, int field, double s) {
    if (o.norm_has[field] && o.norm_sd[field] > 0.0) {
        double mean = o.norm_mean[field], sd = o.norm_sd[field];
        return (s - mean) / sd;
    }
    return 6.0 * (s - 0.5);
}

// src/recommender_similarity.cpp:10-124 profile_similarity(A,B,text_columns_internal).
// fields_out (optional, [3+T]): the float-rounded region/clubs/friends/text similarities, -1 = absent.
float fas(const Oracle& o, int a, int b, float* fields_out) {
    const int T = o.T;
    const int total_possible = PK_NUM_FIXED + T;
    int used = 0;
    double sum_Si = 0.0;
    if (fields_out) for (int j = 0; j < 3 + T; ++j) fields_out[j] = -1.0f;

    if (o.pub[a] >= 0 && o.pub[b] >= 0) {                                     // :38-43
        double s = (o.pub[a] == o.pub[b]) ? 1.0 : 0.0;
        sum_Si += sigmoid(z_of(o, 0, s)); ++used;
    }
    if (o.gender[a] >= 0 && o.gender[b] >= 0) {                               // :45-50
        double s = (o.gender[a] == o.gender[b]) ? 1.0 : 0.0;
        sum_Si += sigmoid(z_of(o, 1, s)); ++used;
    }
    if (o.comp[a] > 0 && o.comp[b] > 0) {                                     // :52-59
        int amin = std::min(o.comp[a], o.comp[b]), amax = std::max(o.comp[a], o.comp[b]);
        double s = (amax > 0) ? ((double)amin / (double)amax) : 0.0;
        sum_Si += sigmoid(z_of(o, 2, s)); ++used;
    }
    if (o.age[a] > 0 && o.age[b] > 0) {                                       // :61-68
        int amin = std::min(o.age[a], o.age[b]), amax = std::max(o.age[a], o.age[b]);
        double s = (amax > 0) ? ((double)amin / (double)amax) : 0.0;
        sum_Si += sigmoid(z_of(o, 3, s)); ++used;
    }
    const int32_t *ra = &o.region[3 * (int64_t)a], *rb = &o.region[3 * (int64_t)b];
    bool nea = ra[0] >= 0 || ra[1] >= 0 || ra[2] >= 0, neb = rb[0] >= 0 || rb[1] >= 0 || rb[2] >= 0;
    if (nea && neb) {                                                         // :70-77
        float f = region_similarity(ra, rb);
        if (fields_out) fields_out[0] = f;
        sum_Si += sigmoid(z_of(o, 4, (double)f)); ++used;
    }
    {
        int64_t a0 = o.clubs_ptr[a], a1 = o.clubs_ptr[a + 1], b0 = o.clubs_ptr[b], b1 = o.clubs_ptr[b + 1];
        if (a1 > a0 && b1 > b0) {                                             // :79-84
            float f = set_similarity(&o.clubs_sorted[a0], a1 - a0, &o.clubs[b0], b1 - b0, nullptr);
            if (fields_out) fields_out[1] = f;
            sum_Si += sigmoid(z_of(o, 5, (double)f)); ++used;
        }
    }
    {
        int64_t a0 = o.friends_ptr[a], a1 = o.friends_ptr[a + 1], b0 = o.friends_ptr[b], b1 = o.friends_ptr[b + 1];
        if (a1 > a0 && b1 > b0) {                                             // :86-91
            float f = set_similarity(&o.friends_sorted[a0], a1 - a0, &o.friends[b0], b1 - b0, nullptr);
            if (fields_out) fields_out[2] = f;
            sum_Si += sigmoid(z_of(o, 6, (double)f)); ++used;
        }
    }
    for (int t = 0; t < T; ++t) {                                             // :93-114
        int64_t a0 = o.tok_ptr[(int64_t)a * T + t], a1 = o.tok_ptr[(int64_t)a * T + t + 1];
        int64_t b0 = o.tok_ptr[(int64_t)b * T + t], b1 = o.tok_ptr[(int64_t)b * T + t + 1];
        if (a1 == a0 || b1 == b0) continue;
        float f;
        if (o.idf[t].present)
            f = tfidf_cosine(&o.tok_id[a0], &o.tok_tf[a0], a1 - a0, &o.tok_id[b0], &o.tok_tf[b0], b1 - b0, o.idf[t]);
        else
            f = raw_cosine(&o.tok_id[a0], &o.tok_tf[a0], a1 - a0, &o.tok_id[b0], &o.tok_tf[b0], b1 - b0);
        if (fields_out) fields_out[3 + t] = f;
        sum_Si += sigmoid(z_of(o, PK_NUM_FIXED + t, (double)f)); ++used;
    }
    if (used == 0) return 0.0f;                                               // :116
    double S = sum_Si / (double)used;
    double F = (double)used / (double)total_possible;
    if (S <= 0.0 && F <= 0.0) return 0.0f;
    double v = (2.0 * S * F) / (S + F);                                       // :122
    return (float)v;
}

// src/recommender_graph.cpp:97-100 comparator
inline bool pair_less(const std::pair<int, float>& A, const std::pair<int, float>& B) {
    if (A.second == B.second) return A.first < B.first;
    return A.second > B.second;
}

typedef std::vector<std::pair<int, float>> Result;

const int32_t* adj_begin(const Oracle& o, int r) { return &o.adj[0] + o.adj_ptr[r]; }
const int32_t* adj_end(const Oracle& o, int r) { return &o.adj[0] + o.adj_ptr[r + 1]; }

// src/recommender_graph.cpp:10-31 gather_candidates_local
void gather_candidates(const Oracle& o, int user, std::vector<int>& out, int limit) {
    out.clear();
    int r = o.adj_row(user);
    if (r < 0) return;
    std::unordered_map<int, char> seen;
    for (const int32_t* pf = adj_begin(o, r); pf != adj_end(o, r); ++pf) {
        int f = *pf;
        if (f == user) continue;
        if (seen.emplace(f, 1).second) out.push_back(f);
        if ((int)out.size() >= limit) return;
        int r2 = o.adj_row(f);
        if (r2 < 0) continue;
        for (const int32_t* pff = adj_begin(o, r2); pff != adj_end(o, r2); ++pff) {
            int ff = *pff;
            if (ff == user) continue;
            if (seen.emplace(ff, 1).second) {
                out.push_back(ff);
                if ((int)out.size() >= limit) return;
            }
        }
    }
}

void finish(Result& out, int topk) {
    std::sort(out.begin(), out.end(), pair_less);
    if ((int)out.size() > topk) out.resize((size_t)std::max(topk, 0));
}

// src/recommender_graph.cpp:33-57,97-103 recommend_graph_registration (== recommend_by_interest :224-227)
Result recommend_graph(const Oracle& o, int user, int topk, int limit) {
    Result out;
    int q = o.row(user);
    if (q < 0) return out;
    std::vector<int> cands;
    gather_candidates(o, user, cands, limit);
    std::unordered_map<int, char> existing;
    int r = o.adj_row(user);
    if (r >= 0) for (const int32_t* p = adj_begin(o, r); p != adj_end(o, r); ++p) existing[*p] = 1;
    existing[user] = 1;
    for (int c : cands) {
        if (existing.count(c)) continue;
        int rc = o.row(c);
        if (rc < 0) continue;
        out.emplace_back(c, fas(o, q, rc, nullptr));
    }
    finish(out, topk);
    return out;
}

// src/recommender_graph.cpp:114-125 candidate expansion of recommend_collaborative: ordered-dedup
// union of the friends' rows minus the user; existing friends are NOT excluded; the limit check
// runs after every visited neighbour (so limit <= 0 still admits one candidate, as in the reference)
void collab_candidates(const Oracle& o, int user, const std::vector<int>& friends, std::vector<int>& cands,
                       int limit) {
    cands.clear();
    std::unordered_map<int, char> cset;
    for (int f : friends) {
        int r2 = o.adj_row(f);
        if (r2 < 0) continue;
        for (const int32_t* p = adj_begin(o, r2); p != adj_end(o, r2); ++p) {
            int fof = *p;
            if (fof == user) continue;
            if (cset.emplace(fof, 1).second) cands.push_back(fof);
            if ((int)cands.size() >= limit) break;
        }
        if ((int)cands.size() >= limit) break;
    }
}

// src/recommender_graph.cpp:105-136,167-180,213-222 recommend_collaborative (profiles branch)
Result recommend_collab(const Oracle& o, int user, int topk, int limit) {
    Result out;
    std::vector<int> friends;
    int r = o.adj_row(user);
    if (r >= 0) friends.assign(adj_begin(o, r), adj_end(o, r));
    std::vector<int> cands;
    collab_candidates(o, user, friends, cands, limit);                        // :114-125
    int q = o.row(user);
    if (q < 0) return out;                                                    // :130
    std::unordered_map<int, float> sim_u_f;
    for (int f : friends) {                                                   // :132-136
        int rf = o.row(f);
        if (rf < 0) continue;
        sim_u_f[f] = fas(o, q, rf, nullptr);
    }
    for (int cand : cands) {                                                  // :167-180
        if (cand == user) continue;
        double score = 0.0;
        int rc = o.row(cand);
        if (rc < 0) continue;
        for (int f : friends) {
            auto it = sim_u_f.find(f);
            if (it == sim_u_f.end()) continue;
            int rf = o.row(f);
            if (rf < 0) continue;
            double s_f_fof = fas(o, rf, rc, nullptr);
            score += (double)it->second * s_f_fof;
        }
        out.emplace_back(cand, (float)score);                                 // :213
    }
    finish(out, topk);
    return out;
}

// src/recommender_clubs.cpp:10-73 recommend_clubs_collab (candidate_limit ignored by the reference)
Result recommend_clubs(const Oracle& o, int user, int topk) {
    Result out;
    int q = o.row(user);
    if (q < 0) return out;
    std::vector<int> friends;
    int r = o.adj_row(user);
    if (r >= 0) friends.assign(adj_begin(o, r), adj_end(o, r));
    std::unordered_map<int, float> sim_u_f;
    for (int f : friends) {                                                   // :23-28
        int rf = o.row(f);
        if (rf < 0) continue;
        sim_u_f[f] = fas(o, q, rf, nullptr);
    }
    std::unordered_map<int, double> club_scores;
    std::vector<int> order;  // first-touch order, only to make the emit loop deterministic
    std::unordered_map<int, char> user_clubs;
    for (int64_t i = o.clubs_ptr[q]; i < o.clubs_ptr[q + 1]; ++i) user_clubs[(int)o.clubs[i]] = 1;
    auto add = [&](int cid, double v) {
        auto it = club_scores.find(cid);
        if (it == club_scores.end()) { club_scores[cid] = 0.0 + v; order.push_back(cid); }
        else it->second += v;
    };
    for (int f : friends) {                                                   // :34-43
        int rf = o.row(f);
        if (rf < 0) continue;
        double w = sim_u_f.count(f) ? sim_u_f.at(f) : 0.0;
        if (w <= 0.0) continue;
        for (int64_t i = o.clubs_ptr[rf]; i < o.clubs_ptr[rf + 1]; ++i) {
            int cid = (int)o.clubs[i];
            if (user_clubs.count(cid)) continue;
            add(cid, w);
        }
    }
    for (int f : friends) {                                                   // :45-64
        int r2 = o.adj_row(f);
        if (r2 < 0) continue;
        int rf = o.row(f);
        if (rf < 0) continue;
        double wuf = sim_u_f.count(f) ? sim_u_f.at(f) : 0.0;
        if (wuf <= 0.0) continue;
        for (const int32_t* p = adj_begin(o, r2); p != adj_end(o, r2); ++p) {
            int fof = *p;
            if (fof == user) continue;
            int rfof = o.row(fof);
            if (rfof < 0) continue;
            double s_f_fof = fas(o, rf, rfof, nullptr);
            if (s_f_fof <= 0.0) continue;
            double contrib = wuf * s_f_fof;
            for (int64_t i = o.clubs_ptr[rfof]; i < o.clubs_ptr[rfof + 1]; ++i) {
                int cid = (int)o.clubs[i];
                if (user_clubs.count(cid)) continue;
                add(cid, contrib);
            }
        }
    }
    for (int cid : order) out.emplace_back(cid, (float)club_scores[cid]);     // :66
    finish(out, topk);
    return out;
}

void emit(const Result& v, int k, int32_t* out_id, float* out_score, int32_t* out_cnt, int64_t q) {
    int c = (int)std::min<size_t>(v.size(), (size_t)std::max(k, 0));
    out_cnt[q] = c;
    for (int i = 0; i < c; ++i) { out_id[q * k + i] = v[i].first; out_score[q * k + i] = v[i].second; }
    for (int i = c; i < k; ++i) { out_id[q * k + i] = -1; out_score[q * k + i] = 0.0f; }
}

template <class V, class P>
void copy_in(V& v, const P* p, int64_t n) { v.assign(p, p + n); }

}  // namespace

extern "C" {

void* orc_create(const pk_world* w) {
    auto* o = new Oracle();
    const int n = w->n_users, T = w->n_text_cols, m = w->n_adj_rows;
    o->n = n; o->T = T;
    copy_in(o->uid, w->uid, n); copy_in(o->pub, w->pub, n); copy_in(o->comp, w->comp, n);
    copy_in(o->gender, w->gender, n); copy_in(o->age, w->age, n); copy_in(o->region, w->region, 3LL * n);
    copy_in(o->clubs_ptr, w->clubs_ptr, n + 1); copy_in(o->clubs, w->clubs, w->clubs_ptr[n]);
    copy_in(o->friends_ptr, w->friends_ptr, n + 1); copy_in(o->friends, w->friends, w->friends_ptr[n]);
    copy_in(o->tok_ptr, w->tok_ptr, (int64_t)n * T + 1);
    copy_in(o->tok_id, w->tok_id, w->tok_ptr[(int64_t)n * T]);
    copy_in(o->tok_tf, w->tok_tf, w->tok_ptr[(int64_t)n * T]);
    copy_in(o->adj_uid, w->adj_uid, m); copy_in(o->adj_ptr, w->adj_ptr, m + 1); copy_in(o->adj, w->adj, w->adj_ptr[m]);
    copy_in(o->norm_mean, w->norm_mean, PK_NUM_FIXED + T); copy_in(o->norm_sd, w->norm_sd, PK_NUM_FIXED + T);
    copy_in(o->norm_has, w->norm_has, PK_NUM_FIXED + T);
    o->adj.push_back(0);  // keep &adj[0] valid for empty graphs
    o->row_of.reserve(n);
    for (int i = 0; i < n; ++i) o->row_of[o->uid[i]] = i;
    o->adj_row_of.reserve(m);
    for (int r = 0; r < m; ++r) o->adj_row_of[o->adj_uid[r]] = r;
    // cells sorted by token id (the reference's cells are hash maps: order-free)
    for (int64_t c = 0; c < (int64_t)n * T; ++c) {
        int64_t b = o->tok_ptr[c], e = o->tok_ptr[c + 1];
        if (e - b > 1 && !std::is_sorted(&o->tok_id[b], &o->tok_id[b] + (e - b))) {
            std::vector<std::pair<int32_t, int32_t>> tmp;
            for (int64_t j = b; j < e; ++j) tmp.emplace_back(o->tok_id[j], o->tok_tf[j]);
            std::sort(tmp.begin(), tmp.end());
            for (int64_t j = b; j < e; ++j) { o->tok_id[j] = tmp[j - b].first; o->tok_tf[j] = tmp[j - b].second; }
        }
    }
    o->clubs_sorted = o->clubs; o->friends_sorted = o->friends;
    o->clubs_sorted.push_back(0); o->friends_sorted.push_back(0); o->clubs.push_back(0); o->friends.push_back(0);
    for (int i = 0; i < n; ++i) {
        std::sort(o->clubs_sorted.begin() + o->clubs_ptr[i], o->clubs_sorted.begin() + o->clubs_ptr[i + 1]);
        std::sort(o->friends_sorted.begin() + o->friends_ptr[i], o->friends_sorted.begin() + o->friends_ptr[i + 1]);
    }
    o->tok_id.push_back(0); o->tok_tf.push_back(0);
    o->idf.resize(T);
    return o;
}

void orc_destroy(void* h) { delete (Oracle*)h; }

// src/recommender.cpp:43-66 compute_idf_from_profiles (float logf); double_log=1 follows
// src/tfidf_index.cpp:7-30 TFIDFIndex::build ((float)log(1.0 + N/(1.0+df)) in double).
int orc_compute_idf(void* h, int double_log) {
    auto* o = (Oracle*)h;
    const int n = o->n, T = o->T;
    size_t total_users = (size_t)n;
    for (int t = 0; t < T; ++t) {
        std::vector<int32_t> all;
        for (int i = 0; i < n; ++i)
            for (int64_t j = o->tok_ptr[(int64_t)i * T + t]; j < o->tok_ptr[(int64_t)i * T + t + 1]; ++j)
                all.push_back(o->tok_id[j]);
        std::sort(all.begin(), all.end());
        IdfCol c;
        c.present = true;
        for (size_t i = 0; i < all.size();) {
            size_t j = i;
            while (j < all.size() && all[j] == all[i]) ++j;
            int df = (int)(j - i);
            float idf = double_log ? (float)std::log(1.0 + (double)(int)total_users / (1.0 + (double)df))
                                   : logf(1.0f + (float)total_users / (1.0f + (float)df));
            c.token.push_back(all[i]); c.idf.push_back(idf); c.df.push_back(df);
            i = j;
        }
        o->idf[t] = std::move(c);
    }
    return 0;
}

int orc_erase_idf_column(void* h, int col) { ((Oracle*)h)->idf[col] = IdfCol(); return 0; }

int orc_erase_idf_tokens(void* h, int col, int n, const int32_t* tokens) {
    auto& c = ((Oracle*)h)->idf[col];
    if (!c.present) return -1;
    for (int i = 0; i < n; ++i) {
        auto it = std::lower_bound(c.token.begin(), c.token.end(), tokens[i]);
        if (it == c.token.end() || *it != tokens[i]) continue;
        size_t k = it - c.token.begin();
        c.token.erase(c.token.begin() + k); c.idf.erase(c.idf.begin() + k); c.df.erase(c.df.begin() + k);
    }
    return 0;
}

int orc_idf_lookup(void* h, int col, int n, const int32_t* tokens, float* out, uint8_t* has, int32_t* df) {
    const auto& c = ((Oracle*)h)->idf[col];
    for (int i = 0; i < n; ++i) {
        has[i] = 0; out[i] = 0.0f; if (df) df[i] = 0;
        if (!c.present) continue;
        auto it = std::lower_bound(c.token.begin(), c.token.end(), tokens[i]);
        if (it == c.token.end() || *it != tokens[i]) continue;
        size_t k = it - c.token.begin();
        has[i] = 1; out[i] = c.idf[k]; if (df) df[i] = c.df[k];
    }
    return 0;
}

int64_t orc_idf_size(void* h, int col) {
    const auto& c = ((Oracle*)h)->idf[col];
    return c.present ? (int64_t)c.token.size() : -1;
}

int orc_pair_similarity(void* h, int64_t n, const int32_t* a, const int32_t* b, float* out, int nthreads) {
    auto* o = (Oracle*)h;
    parallel_for(n, nthreads, [&](int64_t i) {
        int ra = o->row(a[i]), rb = o->row(b[i]);
        out[i] = (ra < 0 || rb < 0) ? -1.0f : fas(*o, ra, rb, nullptr);
    });
    return 0;
}

int orc_pair_fields(void* h, int64_t n, const int32_t* a, const int32_t* b, float* out, int nthreads) {
    auto* o = (Oracle*)h;
    const int T = o->T;
    parallel_for(n, nthreads, [&](int64_t i) {
        int ra = o->row(a[i]), rb = o->row(b[i]);
        float* f = out + i * (3 + T);
        if (ra < 0 || rb < 0) { for (int j = 0; j < 3 + T; ++j) f[j] = -1.0f; return; }
        fas(*o, ra, rb, f);
    });
    return 0;
}

// set-overlap counts (integer, bit-exact contract): which = 0 clubs, 1 friends
int orc_pair_overlap(void* h, int64_t n, const int32_t* a, const int32_t* b, int which, int32_t* out) {
    auto* o = (Oracle*)h;
    for (int64_t i = 0; i < n; ++i) {
        int ra = o->row(a[i]), rb = o->row(b[i]);
        out[i] = -1;
        if (ra < 0 || rb < 0) continue;
        const auto& ptr = which ? o->friends_ptr : o->clubs_ptr;
        const auto& srt = which ? o->friends_sorted : o->clubs_sorted;
        const auto& raw = which ? o->friends : o->clubs;
        int inter = 0;
        set_similarity(&srt[ptr[ra]], ptr[ra + 1] - ptr[ra], &raw[ptr[rb]], ptr[rb + 1] - ptr[rb], &inter);
        out[i] = inter;
    }
    return 0;
}

int orc_recommend(void* h, int mode, int nq, const int32_t* q, int k, int limit, int32_t* out_id,
                  float* out_score, int32_t* out_cnt, int nthreads) {
    auto* o = (Oracle*)h;
    parallel_for(nq, nthreads, [&](int64_t i) {
        Result v;
        if (mode == PK_MODE_GRAPH) v = recommend_graph(*o, q[i], k, limit);
        else if (mode == PK_MODE_COLLAB) v = recommend_collab(*o, q[i], k, limit);
        else if (mode == PK_MODE_CLUBS) v = recommend_clubs(*o, q[i], k);
        emit(v, k, out_id, out_score, out_cnt, i);
    });
    return 0;
}

// candidate id lists (integer, bit-exact contract).  mode GRAPH: gather_candidates_local output
// (before the existing/profile filter); mode COLLAB: the ordered-dedup FoF expansion (:116-125).
int orc_candidates(void* h, int mode, int user, int limit, int32_t* out, int32_t cap) {
    auto* o = (Oracle*)h;
    std::vector<int> c;
    if (mode == PK_MODE_GRAPH) gather_candidates(*o, user, c, limit);
    else {
        std::vector<int> friends;
        int r = o->adj_row(user);
        if (r >= 0) friends.assign(adj_begin(*o, r), adj_end(*o, r));
        collab_candidates(*o, user, friends, c, limit);
    }
    int nout = (int)std::min<size_t>(c.size(), (size_t)cap);
    for (int i = 0; i < nout; ++i) out[i] = c[i];
    return (int)c.size();
}

// Full-scan extension: every profile is a candidate (minus the exclusion set), reference comparator.
int orc_fullscan(void* h, int nq, const int32_t* q, int k, int exclude_mode, int32_t* out_id,
                 float* out_score, int32_t* out_cnt, int nthreads) {
    auto* o = (Oracle*)h;
    parallel_for(nq, nthreads, [&](int64_t i) {
        Result v;
        int rq = o->row(q[i]);
        if (rq >= 0) {
            std::unordered_map<int, char> ex;
            if (exclude_mode >= PK_EXCLUDE_SELF) ex[q[i]] = 1;
            if (exclude_mode >= PK_EXCLUDE_SELF_AND_ADJ) {
                int r = o->adj_row(q[i]);
                if (r >= 0) for (const int32_t* p = adj_begin(*o, r); p != adj_end(*o, r); ++p) ex[*p] = 1;
            }
            v.reserve(o->n);
            for (int c = 0; c < o->n; ++c) {
                if (!ex.empty() && ex.count(o->uid[c])) continue;
                v.emplace_back(o->uid[c], fas(*o, rq, c, nullptr));
            }
            size_t kk = std::min<size_t>(v.size(), (size_t)std::max(k, 0));
            std::partial_sort(v.begin(), v.begin() + kk, v.end(), pair_less);
            v.resize(kk);
        }
        emit(v, k, out_id, out_score, out_cnt, i);
    });
    return 0;
}

// all candidate scores of one query (used by the tests to explain any top-k id mismatch)
int orc_scores_of_query(void* h, int32_t quid, float* out /*[n]*/, int nthreads) {
    auto* o = (Oracle*)h;
    int rq = o->row(quid);
    if (rq < 0) return -1;
    parallel_for(o->n, nthreads, [&](int64_t c) { out[c] = fas(*o, rq, (int)c, nullptr); });
    return 0;
}

}  // extern "C"
