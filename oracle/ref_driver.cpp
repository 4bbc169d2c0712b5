// ref_driver.cpp -- TEST INFRASTRUCTURE.  Thin C ABI around the UNMODIFIED compiled reference
// (pymlex/recommendation-system-pokec).  It is compiled together with the reference's own sources
// *where they lie* under /root/reference (see oracle/build_oracle.py); output goes to oracle/_ref/.
// Nothing here re-implements any arithmetic: it fills the reference's own containers from the flat
// pk_world arrays and calls the reference's public Recommender API (include/recommender.h:17-71).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
// the resulting library.
#include "recommender.h"      // reference header, from /root/reference/include
#include "user_profile.h"
#ifndef PK_DROPIN_BUILD
#include "utils.h"
#endif
#include "tfidf_index.h"
#include "test.h"             // reference header: run_friends_holdout_test (src/test.cpp:13-105)

#include "pokec_world.h"

#include <algorithm>
#include <atomic>
#include <cstring>
#include <random>
#include <unordered_set>
#include <string>
#include <thread>
#include <vector>

using namespace std;

// The reference declares `friend struct ::RecommenderInternalSim` (include/recommender.h:15,70) and
// never defines it.  Defining it gives the tests access to the private per-field helpers so that each
// float rounding point can be pinned separately.  Inside the drop-in library the definition is the
// drop-in's own header (one definition per library); the compiled-reference build defines it here.
#ifdef PK_DROPIN_BUILD
#include "recommender_internal_sim.h"
#else
struct RecommenderInternalSim {
    static float set_sim(const vector<uint32_t>& a, const vector<uint32_t>& b) {
        return Recommender::vec_set_similarity(a, b);
    }
    static float region_sim(const array<int, 3>& a, const array<int, 3>& b) {
        return Recommender::region_similarity_local(a, b);
    }
    static float raw_cos(const unordered_map<int, int>& a, const unordered_map<int, int>& b) {
        return Recommender::cosine_counts_maps_local(a, b);
    }
    static float tfidf_cos(const Recommender& r, const unordered_map<int, int>& a,
                           const unordered_map<int, int>& b, const unordered_map<int, float>& idf) {
        return r.tfidf_cosine_for_column(a, b, idf);
    }
};
#endif

namespace {

struct RefHandle {
    unordered_map<int, UserProfile> profiles;
    unordered_map<int, vector<int>> adj;
    vector<string> cols;
    Recommender* rec = nullptr;
    ~RefHandle() { delete rec; }
};

template <class F>
void parallel_for(int64_t n, int nthreads, F f) {
    if (nthreads <= 1 || n <= 1) {
        for (int64_t i = 0; i < n; ++i) f(i);
        return;
    }
    atomic<int64_t> next(0);
    vector<thread> th;
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

inline bool pair_less(const pair<int, float>& A, const pair<int, float>& B) {
    // verbatim semantics of the reference comparator (src/recommender_graph.cpp:97-100)
    if (A.second == B.second) return A.first < B.first;
    return A.second > B.second;
}

void emit(const vector<pair<int, float>>& v, int k, int32_t* out_id, float* out_score, int32_t* out_cnt,
          int64_t q) {
    int c = (int)min<size_t>(v.size(), (size_t)max(k, 0));
    out_cnt[q] = c;
    for (int i = 0; i < c; ++i) {
        out_id[q * k + i] = v[i].first;
        out_score[q * k + i] = v[i].second;
    }
    for (int i = c; i < k; ++i) {
        out_id[q * k + i] = -1;
        out_score[q * k + i] = 0.0f;
    }
}

}  // namespace

extern "C" {

void* ref_create(const pk_world* w, const char* const* col_names) {
    auto* h = new RefHandle();
    const int n = w->n_users, T = w->n_text_cols;
    h->profiles.reserve((size_t)n);
    for (int i = 0; i < n; ++i) {
        UserProfile p;
        p.user_id = w->uid[i];
        p.public_flag = w->pub[i];
        p.completion_percentage = w->comp[i];
        p.gender = w->gender[i];
        p.age = w->age[i];
        p.clubs.assign(w->clubs + w->clubs_ptr[i], w->clubs + w->clubs_ptr[i + 1]);
        p.friends.assign(w->friends + w->friends_ptr[i], w->friends + w->friends_ptr[i + 1]);
        p.region_parts = {w->region[3 * i], w->region[3 * i + 1], w->region[3 * i + 2]};
        p.token_cols.resize((size_t)T);
        for (int t = 0; t < T; ++t) {
            int64_t b = w->tok_ptr[(int64_t)i * T + t], e = w->tok_ptr[(int64_t)i * T + t + 1];
            for (int64_t j = b; j < e; ++j) p.token_cols[t][w->tok_id[j]] = w->tok_tf[j];
        }
        h->profiles.emplace(p.user_id, std::move(p));
    }
    h->adj.reserve((size_t)w->n_adj_rows);
    for (int r = 0; r < w->n_adj_rows; ++r)
        h->adj[w->adj_uid[r]].assign(w->adj + w->adj_ptr[r], w->adj + w->adj_ptr[r + 1]);
    for (int t = 0; t < T; ++t) h->cols.emplace_back(col_names[t]);

    h->rec = new Recommender(&h->profiles, &h->adj);
    static const char* fixed[PK_NUM_FIXED] = {"public", "gender", "completion", "age",
                                               "region", "clubs",  "friends"};
    unordered_map<string, pair<float, float>> fm, cm;
    for (int f = 0; f < PK_NUM_FIXED; ++f)
        if (w->norm_has && w->norm_has[f]) fm[fixed[f]] = {w->norm_mean[f], w->norm_sd[f]};
    for (int t = 0; t < T; ++t)
        if (w->norm_has && w->norm_has[PK_NUM_FIXED + t])
            cm[h->cols[t]] = {w->norm_mean[PK_NUM_FIXED + t], w->norm_sd[PK_NUM_FIXED + t]};
    h->rec->set_field_normalizers(fm);
    h->rec->set_column_normalizers(cm);
    h->rec->set_text_columns(h->cols);
    return h;
}

void ref_destroy(void* hv) { delete (RefHandle*)hv; }

// compute_idf_from_profiles (src/recommender.cpp:43-66)
int ref_compute_idf(void* hv) {
    auto* h = (RefHandle*)hv;
    h->rec->compute_idf_from_profiles(h->cols);
    return 0;
}

// drop the idf entry of one column -> that column falls back to raw-count cosine
// (src/recommender_similarity.cpp:99-104)
int ref_erase_idf_column(void* hv, int col) {
    auto* h = (RefHandle*)hv;
    h->rec->idf_per_col.erase(h->cols[col]);
    return 0;
}

// remove single tokens from a column's idf map -> they get idf 1.0f (src/recommender.cpp:78)
int ref_erase_idf_tokens(void* hv, int col, int n, const int32_t* tokens) {
    auto* h = (RefHandle*)hv;
    auto it = h->rec->idf_per_col.find(h->cols[col]);
    if (it == h->rec->idf_per_col.end()) return -1;
    for (int i = 0; i < n; ++i) it->second.erase(tokens[i]);
    return 0;
}

// idf lookup; has[i]=0 when the token (or the column) has no entry
int ref_idf_lookup(void* hv, int col, int n, const int32_t* tokens, float* out, uint8_t* has) {
    auto* h = (RefHandle*)hv;
    auto it = h->rec->idf_per_col.find(h->cols[col]);
    for (int i = 0; i < n; ++i) {
        has[i] = 0;
        out[i] = 0.0f;
        if (it == h->rec->idf_per_col.end()) continue;
        auto jt = it->second.find(tokens[i]);
        if (jt == it->second.end()) continue;
        has[i] = 1;
        out[i] = jt->second;
    }
    return 0;
}

int64_t ref_idf_size(void* hv, int col) {
    auto* h = (RefHandle*)hv;
    auto it = h->rec->idf_per_col.find(h->cols[col]);
    return it == h->rec->idf_per_col.end() ? -1 : (int64_t)it->second.size();
}

// TFIDFIndex::build (src/tfidf_index.cpp:7-30): double-log idf variant
int ref_tfidf_index_lookup(void* hv, int col, int n, const int32_t* tokens, float* out, uint8_t* has) {
    auto* h = (RefHandle*)hv;
    TFIDFIndex idx;
    idx.build(h->profiles, h->cols);
    auto it = idx.idf_per_col.find(h->cols[col]);
    for (int i = 0; i < n; ++i) {
        has[i] = 0;
        out[i] = 0.0f;
        if (it == idx.idf_per_col.end()) continue;
        auto jt = it->second.find(tokens[i]);
        if (jt == it->second.end()) continue;
        has[i] = 1;
        out[i] = jt->second;
    }
    return 0;
}

// Recommender::profile_similarity (src/recommender_similarity.cpp:10-128)
int ref_pair_similarity(void* hv, int64_t n, const int32_t* a, const int32_t* b, float* out, int nthreads) {
    auto* h = (RefHandle*)hv;
    parallel_for(n, nthreads, [&](int64_t i) {
        auto ia = h->profiles.find(a[i]);
        auto ib = h->profiles.find(b[i]);
        out[i] = (ia == h->profiles.end() || ib == h->profiles.end())
                     ? -1.0f
                     : h->rec->profile_similarity(ia->second, ib->second);
    });
    return 0;
}

// per-field raw similarities (float rounding points of SURVEY App. A item 3).  out is [n][3+T]:
// region, clubs, friends, text columns.  Fields absent on either side are reported as -1.
int ref_pair_fields(void* hv, int64_t n, const int32_t* a, const int32_t* b, float* out, int nthreads) {
    auto* h = (RefHandle*)hv;
    const int T = (int)h->cols.size();
    parallel_for(n, nthreads, [&](int64_t i) {
        float* o = out + i * (3 + T);
        for (int j = 0; j < 3 + T; ++j) o[j] = -1.0f;
        auto ia = h->profiles.find(a[i]);
        auto ib = h->profiles.find(b[i]);
        if (ia == h->profiles.end() || ib == h->profiles.end()) return;
        const UserProfile &A = ia->second, &B = ib->second;
        auto any = [](const array<int, 3>& r) { return r[0] >= 0 || r[1] >= 0 || r[2] >= 0; };
        if (any(A.region_parts) && any(B.region_parts))
            o[0] = RecommenderInternalSim::region_sim(A.region_parts, B.region_parts);
        if (!A.clubs.empty() && !B.clubs.empty()) o[1] = RecommenderInternalSim::set_sim(A.clubs, B.clubs);
        if (!A.friends.empty() && !B.friends.empty())
            o[2] = RecommenderInternalSim::set_sim(A.friends, B.friends);
        for (int t = 0; t < T; ++t) {
            if (A.token_cols[t].empty() || B.token_cols[t].empty()) continue;
            auto it = h->rec->idf_per_col.find(h->cols[t]);
            o[3 + t] = (it != h->rec->idf_per_col.end())
                           ? RecommenderInternalSim::tfidf_cos(*h->rec, A.token_cols[t], B.token_cols[t], it->second)
                           : RecommenderInternalSim::raw_cos(A.token_cols[t], B.token_cols[t]);
        }
    });
    return 0;
}

// the four recommend_* calls (src/recommender_graph.cpp:33-227, src/recommender_clubs.cpp:10-73)
int ref_recommend(void* hv, int mode, int nq, const int32_t* q, int k, int limit, int32_t* out_id,
                  float* out_score, int32_t* out_cnt, int nthreads) {
    auto* h = (RefHandle*)hv;
    parallel_for(nq, nthreads, [&](int64_t i) {
        vector<pair<int, float>> v;
        switch (mode) {
            case PK_MODE_GRAPH: v = h->rec->recommend_by_interest(q[i], k, limit); break;
            case PK_MODE_COLLAB: v = h->rec->recommend_collaborative(q[i], k, limit); break;
            case PK_MODE_CLUBS: v = h->rec->recommend_clubs_collab(q[i], k, limit); break;
            default: break;
        }
        emit(v, k, out_id, out_score, out_cnt, i);
    });
    return 0;
}

// Full-scan extension: candidate set = every profile, scored with the
// reference's own profile_similarity and ordered by the reference's own comparator.
int ref_fullscan(void* hv, int nq, const int32_t* q, int k, int exclude_mode, int32_t* out_id,
                 float* out_score, int32_t* out_cnt, int nthreads) {
    auto* h = (RefHandle*)hv;
    parallel_for(nq, nthreads, [&](int64_t i) {
        vector<pair<int, float>> v;
        auto itq = h->profiles.find(q[i]);
        if (itq != h->profiles.end()) {
            unordered_map<int, char> ex;
            if (exclude_mode >= PK_EXCLUDE_SELF) ex[q[i]] = 1;
            if (exclude_mode >= PK_EXCLUDE_SELF_AND_ADJ) {
                auto ita = h->adj.find(q[i]);
                if (ita != h->adj.end())
                    for (int f : ita->second) ex[f] = 1;
            }
            v.reserve(h->profiles.size());
            for (auto& kv : h->profiles) {
                if (ex.count(kv.first)) continue;
                v.emplace_back(kv.first, h->rec->profile_similarity(itq->second, kv.second));
            }
            size_t kk = min<size_t>(v.size(), (size_t)max(k, 0));
            partial_sort(v.begin(), v.begin() + kk, v.end(), pair_less);
            v.resize(kk);
        }
        emit(v, k, out_id, out_score, out_cnt, i);
    });
    return 0;
}

#ifndef PK_DROPIN_BUILD   // utils.cpp is outside the replaced path; the drop-in build does not link it
// compute_column_normalizers (src/utils.cpp:155-240), seed 12345 inside the reference
int ref_column_normalizers(void* hv, int sample_size, int comps_per_user, float* mean, float* sd) {
    auto* h = (RefHandle*)hv;
    auto m = compute_column_normalizers(h->profiles, h->cols, sample_size, comps_per_user);
    static const char* fixed[PK_NUM_FIXED] = {"public", "gender", "completion", "age",
                                               "region", "clubs",  "friends"};
    for (int f = 0; f < PK_NUM_FIXED; ++f) {
        auto it = m.find(fixed[f]);
        mean[f] = it == m.end() ? 0.f : it->second.first;
        sd[f] = it == m.end() ? 0.f : it->second.second;
    }
    for (size_t t = 0; t < h->cols.size(); ++t) {
        auto it = m.find(h->cols[t]);
        mean[PK_NUM_FIXED + t] = it == m.end() ? 0.f : it->second.first;
        sd[PK_NUM_FIXED + t] = it == m.end() ? 0.f : it->second.second;
    }
    return 0;
}
#endif

// The pair sample compute_column_normalizers draws (src/utils.cpp:163-181): ids in the iteration order of the
// reference's unordered_map, mt19937(12345), uniform_int_distribution over the ids, a == b and repeated pairs
// skipped, at most 10 x the wanted count of attempts.  Restated here (the reference does not expose it) with the
// same libstdc++ so that the device path can be fed the very pairs the reference scores.  Returns the count.
long long ref_normalizer_pairs(void* hv, int sample_size, int comps_per_user, int32_t* a_out, int32_t* b_out, long long cap) {
    auto* h = (RefHandle*)hv;
    if (h->profiles.empty()) return 0;
    vector<int> ids;
    ids.reserve(h->profiles.size());
    for (auto& kv : h->profiles) ids.push_back(kv.first);
    mt19937 rng(12345);
    uniform_int_distribution<size_t> dist(0, ids.size() - 1);
    const size_t total_needed = (size_t)sample_size * (size_t)comps_per_user;
    unordered_set<uint64_t> seen;
    size_t attempts = 0;
    long long n = 0;
    while (seen.size() < total_needed && attempts < total_needed * 10) {
        ++attempts;
        const int a = ids[dist(rng)];
        const int b = ids[dist(rng)];
        if (a == b) continue;
        uint32_t A = (uint32_t)a, B = (uint32_t)b;
        if (A > B) swap(A, B);
        if (!seen.insert(((uint64_t)A << 32) | (uint64_t)B).second) continue;
        if (n < cap) { a_out[n] = a; b_out[n] = b; }
        ++n;
    }
    return n;
}

#ifdef PK_DROPIN_BUILD
}  // extern "C"
#include "recommender_b200_batch.h"
extern "C" {
// SURVEY 8 f2 through the reference's own containers: the handle's Recommender is written with
// pokec_b200::save_world_file, read back with load_world_file into fresh maps, and the two worlds are compared
// member by member.  0 = identical; otherwise the number of the first check that failed.
int ref_world_file_round_trip(void* hv, const char* path) {
    auto* h = (RefHandle*)hv;
    if (!pokec_b200::save_world_file(*h->rec, path)) return 1;
    unordered_map<int, UserProfile> P;
    unordered_map<int, vector<int>> A;
    vector<float> mean, sd;
    vector<uint8_t> has;
    const int T = pokec_b200::load_world_file(path, P, A, &mean, &sd, &has);
    if (T != (int)h->cols.size()) return 2;
    if (P.size() != h->profiles.size() || A.size() != h->adj.size()) return 3;
    for (auto& kv : h->profiles) {
        auto it = P.find(kv.first);
        if (it == P.end()) return 4;
        const UserProfile &a = kv.second, &b = it->second;
        if (b.user_id != kv.first || a.public_flag != b.public_flag || a.completion_percentage != b.completion_percentage ||
            a.gender != b.gender || a.age != b.age || a.region_parts != b.region_parts) return 5;
        if (a.clubs != b.clubs || a.friends != b.friends) return 6;
        if ((int)b.token_cols.size() != T) return 7;
        for (int t = 0; t < T; ++t) {
            static const unordered_map<int, int> none;
            const auto& ca = (size_t)t < a.token_cols.size() ? a.token_cols[t] : none;
            if (ca != b.token_cols[t]) return 8;
        }
    }
    for (auto& kv : h->adj) {
        auto it = A.find(kv.first);
        if (it == A.end() || it->second != kv.second) return 9;
    }
    if ((int)mean.size() != PK_NUM_FIXED + T || sd.size() != mean.size() || has.size() != mean.size()) return 10;
    static const char* fixed[PK_NUM_FIXED] = {"public", "gender", "completion", "age", "region", "clubs", "friends"};
    for (int f = 0; f < PK_NUM_FIXED + T; ++f) {
        const auto& m = f < PK_NUM_FIXED ? h->rec->field_normalizers : h->rec->column_normalizers;
        auto it = m.find(f < PK_NUM_FIXED ? string(fixed[f]) : h->cols[f - PK_NUM_FIXED]);
        if ((it != m.end()) != (has[f] != 0)) return 11;
        if (it != m.end() && (memcmp(&it->second.first, &mean[f], 4) != 0 || memcmp(&it->second.second, &sd[f], 4) != 0)) return 12;
    }
    unordered_map<int, UserProfile> P2;
    unordered_map<int, vector<int>> A2;
    if (pokec_b200::load_world_file(string(path) + ".missing", P2, A2) != -1) return 13;
    return 0;
}
#endif

// The reference's only live "test" (src/main.cpp:147 -> src/test.cpp:13-105), UNMODIFIED and compiled where it lies:
// it builds its own Recommender over a copy of the adjacency, removes a fifth of a user's friends behind it, asks
// recommend_collaborative(uid, hold_k, 1000) and writes one hit ratio per tested user to out_path.  In the drop-in
// library the same caller code runs against the CUDA-backed Recommender.
int ref_holdout_test(void* hv, int sample_size, const char* out_path) {
    auto* h = (RefHandle*)hv;
    run_friends_holdout_test(h->profiles, h->adj, h->cols, *h->rec, sample_size, out_path);
    return 0;
}

// mutate one adjacency row behind the live Recommender (what src/test.cpp:59-73 does)
int ref_set_adj_row(void* hv, int uid, int n, const int32_t* nb) {
    auto* h = (RefHandle*)hv;
    h->adj[uid].assign(nb, nb + n);
    return 0;
}

}  // extern "C"
