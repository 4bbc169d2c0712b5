// recommender_b200.cpp -- the drop-in: every member of the reference's `Recommender`
// (pymlex/recommendation-system-pokec include/recommender.h:17-71, header UNCHANGED and not copied:
// this file is compiled with -I <reference>/include) implemented on top of the C ABI of
// libpokec_b200.so (include/pokec_b200.h).  It replaces the reference translation units
//   src/recommender.cpp  src/recommender_similarity.cpp  src/recommender_graph.cpp  src/recommender_clubs.cpp
// at link time; api_cli.cpp / main.cpp / test.cpp / ui.cpp compile unmodified against it.
//
// What runs here on the host is what the C ABI needs as input: flattening the caller-owned hash maps
// into the pk_world arrays once per Recommender, pushing the (public, caller-mutable) normalizer / idf
// members when they changed, and re-sending the adjacency rows a query touches (the owner may rewrite
// them behind the Recommender, src/test.cpp:59-73).  Every similarity, df count and top-k runs in the
// CUDA library; there is no CPU scoring path -- a CUDA failure is fatal (message on stderr, never stdout:
// stdout is api_cli's JSON channel, src/api_cli.cpp:172-255).
#include "recommender.h"
#include "user_profile.h"

#include "pokec_b200.h"
#include "recommender_b200_batch.h"
#include "recommender_internal_sim.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

using namespace std;

namespace {

const char* kFixedKeys[PK_NUM_FIXED] = {"public", "gender", "completion", "age", "region", "clubs", "friends"};

[[noreturn]] void die(const char* what) {
    fprintf(stderr, "[pokec_b200] fatal: %s: %s\n", what, pk_last_error());
    abort();
}
#define PKCHECK(call) do { if ((call) != PK_OK) die(#call); } while (0)

uint64_t mix(uint64_t h, uint64_t v) {
    h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
    return h;
}
uint64_t hash_str(const string& s) {
    uint64_t h = 1469598103934665603ull;
    for (unsigned char c : s) { h ^= c; h *= 1099511628211ull; }
    return h;
}
uint32_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

// fingerprint of one idf column map: size + a few entries from the front of the iteration order.  The
// members are public and callers assign them wholesale (src/test.cpp:37-40), never token by token.
uint64_t idf_fingerprint(const unordered_map<int, float>& m) {
    uint64_t h = mix(0x1d0f, m.size());
    int seen = 0;
    for (auto it = m.begin(); it != m.end() && seen < 8; ++it, ++seen) h = mix(mix(h, (uint32_t)it->first), fbits(it->second));
    return h;
}
constexpr uint64_t kAbsent = 0xabcdefabcdefull;

// flat arrays behind a pk_world
struct Flat {
    vector<int32_t> uid, pub, comp, gender, age, region, tok_id, tok_tf, adj_uid, adj;
    vector<int64_t> clubs_ptr, friends_ptr, tok_ptr, adj_ptr;
    vector<uint32_t> clubs, friends;
    vector<float> nmean, nsd;
    vector<uint8_t> nhas;
    pk_world w;
    void add_profile(const UserProfile& p, int32_t as_uid, int T) {
        uid.push_back(as_uid);
        pub.push_back(p.public_flag); comp.push_back(p.completion_percentage); gender.push_back(p.gender); age.push_back(p.age);
        for (int i = 0; i < 3; ++i) region.push_back(p.region_parts[i]);
        clubs.insert(clubs.end(), p.clubs.begin(), p.clubs.end());
        clubs_ptr.push_back((int64_t)clubs.size());
        friends.insert(friends.end(), p.friends.begin(), p.friends.end());
        friends_ptr.push_back((int64_t)friends.size());
        for (int t = 0; t < T; ++t) {
            if ((size_t)t < p.token_cols.size())
                for (auto& kv : p.token_cols[t]) { tok_id.push_back(kv.first); tok_tf.push_back(kv.second); }
            tok_ptr.push_back((int64_t)tok_id.size());
        }
    }
    void begin() { clubs_ptr.assign(1, 0); friends_ptr.assign(1, 0); tok_ptr.assign(1, 0); adj_ptr.assign(1, 0); }
    const pk_world* finish(int T) {
        memset(&w, 0, sizeof(w));
        w.n_users = (int32_t)uid.size(); w.n_text_cols = T;
        w.uid = uid.data(); w.pub = pub.data(); w.comp = comp.data(); w.gender = gender.data(); w.age = age.data();
        w.region = region.data(); w.clubs_ptr = clubs_ptr.data(); w.clubs = clubs.data();
        w.friends_ptr = friends_ptr.data(); w.friends = friends.data();
        w.tok_ptr = tok_ptr.data(); w.tok_id = tok_id.data(); w.tok_tf = tok_tf.data();
        w.n_adj_rows = (int32_t)adj_uid.size(); w.adj_uid = adj_uid.data(); w.adj_ptr = adj_ptr.data(); w.adj = adj.data();
        return &w;
    }
};

struct Shim {
    pk_store* store = nullptr;
    pk_store* scratch = nullptr;          // 2-profile store for profile_similarity on arbitrary objects
    const void* profiles_ptr = nullptr;
    const void* adj_ptr = nullptr;
    size_t n_profiles = 0, n_adj = 0;
    int T = -1;
    uint64_t norm_fp = 0;
    vector<uint64_t> idf_fp;              // per column: fingerprint of the table the device holds
    vector<string> names;                 // column names the device tables were pushed under
};

mutex g_mu;
unordered_map<const Recommender*, Shim> g_shims;

int device_from_env() {
    const char* e = getenv("POKEC_B200_DEVICE");
    return e ? atoi(e) : 0;
}

void drop(Shim& s) {
    if (s.store) pk_store_destroy(s.store);
    if (s.scratch) pk_store_destroy(s.scratch);
    s = Shim();
}

}  // namespace

namespace {

// ---- world upload ----------------------------------------------------------------------------------
// rows are sorted by uid so that a row range of the store is a user-id range (the multi-GPU shard key)
void flatten(const Recommender& r, int T, Flat& f) {
    f.begin();
    const auto& P = *r.profiles;
    vector<const UserProfile*> order;
    vector<int> keys;
    order.reserve(P.size());
    vector<pair<int, const UserProfile*>> kv;
    kv.reserve(P.size());
    for (auto& e : P) kv.emplace_back(e.first, &e.second);
    sort(kv.begin(), kv.end(), [](const pair<int, const UserProfile*>& a, const pair<int, const UserProfile*>& b) { return a.first < b.first; });
    for (auto& e : kv) f.add_profile(*e.second, e.first, T);   // the MAP KEY is the profile's identity (recommender_graph.cpp:53)
    if (r.adj_list) {
        vector<pair<int, const vector<int>*>> rows;
        rows.reserve(r.adj_list->size());
        for (auto& e : *r.adj_list) rows.emplace_back(e.first, &e.second);
        sort(rows.begin(), rows.end(), [](const pair<int, const vector<int>*>& a, const pair<int, const vector<int>*>& b) { return a.first < b.first; });
        for (auto& e : rows) {
            f.adj_uid.push_back(e.first);
            f.adj.insert(f.adj.end(), e.second->begin(), e.second->end());
            f.adj_ptr.push_back((int64_t)f.adj.size());
        }
    }
}

void upload_world(Shim& s, const Recommender& r, int T) {
    if (!s.store) PKCHECK(pk_store_create(device_from_env(), &s.store));
    Flat f;
    flatten(r, T, f);
    const auto& P = *r.profiles;
    PKCHECK(pk_store_load_world(s.store, f.finish(T)));
    s.profiles_ptr = r.profiles; s.adj_ptr = r.adj_list;
    s.n_profiles = P.size(); s.n_adj = r.adj_list ? r.adj_list->size() : 0;
    s.T = T;
    s.norm_fp = 0;
    s.idf_fp.assign((size_t)T, kAbsent);      // a freshly loaded store has no idf table for any column
    s.names.assign((size_t)T, string());
}

uint64_t norm_fingerprint(const Recommender& r, const vector<string>& cols) {
    uint64_t h = 7;
    for (int f = 0; f < PK_NUM_FIXED; ++f) {
        auto it = r.field_normalizers.find(kFixedKeys[f]);
        h = (it == r.field_normalizers.end()) ? mix(h, kAbsent) : mix(mix(h, fbits(it->second.first)), fbits(it->second.second));
    }
    for (auto& c : cols) {
        auto it = r.column_normalizers.find(c);
        h = (it == r.column_normalizers.end()) ? mix(h, kAbsent) : mix(mix(h, fbits(it->second.first)), fbits(it->second.second));
    }
    return h;
}

void normalizer_arrays(const Recommender& r, const vector<string>& cols, vector<float>& mean, vector<float>& sd, vector<uint8_t>& has) {
    const int nf = PK_NUM_FIXED + (int)cols.size();
    mean.assign(nf, 0.f); sd.assign(nf, 0.f); has.assign(nf, 0);
    for (int f = 0; f < PK_NUM_FIXED; ++f) {
        auto it = r.field_normalizers.find(kFixedKeys[f]);                 // recommender_similarity.cpp:28-36
        if (it != r.field_normalizers.end()) { mean[f] = it->second.first; sd[f] = it->second.second; has[f] = 1; }
    }
    for (size_t t = 0; t < cols.size(); ++t) {
        auto it = r.column_normalizers.find(cols[t]);                      // :105-111
        if (it != r.column_normalizers.end()) { mean[PK_NUM_FIXED + t] = it->second.first; sd[PK_NUM_FIXED + t] = it->second.second; has[PK_NUM_FIXED + t] = 1; }
    }
}

void push_normalizers(pk_store* st, const Recommender& r, const vector<string>& cols) {
    vector<float> mean, sd;
    vector<uint8_t> has;
    normalizer_arrays(r, cols, mean, sd, has);
    PKCHECK(pk_set_normalizers(st, mean.data(), sd.data(), has.data(), (int)mean.size()));
}

// Brings the device store in line with the Recommender's current public state.  `cols` is the column
// list the call scores with (text_columns_internal, or the list passed to compute_idf_from_profiles).
Shim& sync_state(const Recommender& r, const vector<string>& cols) {
    Shim& s = g_shims[&r];
    const int T = (int)cols.size();
    const bool stale = !s.store || s.profiles_ptr != r.profiles || s.adj_ptr != r.adj_list || s.T != T ||
                       s.n_profiles != r.profiles->size() || s.n_adj != (r.adj_list ? r.adj_list->size() : 0);
    if (stale) upload_world(s, r, T);
    const uint64_t nfp = mix(norm_fingerprint(r, cols), 1);
    if (nfp != s.norm_fp) { push_normalizers(s.store, r, cols); s.norm_fp = nfp; }
    vector<int32_t> tok;
    vector<float> val;
    for (int t = 0; t < T; ++t) {
        auto it = r.idf_per_col.find(cols[t]);
        const uint64_t fp = (it == r.idf_per_col.end()) ? kAbsent : mix(idf_fingerprint(it->second), hash_str(cols[t]));
        if (fp == s.idf_fp[t]) continue;
        if (it == r.idf_per_col.end()) {
            PKCHECK(pk_set_idf(s.store, t, nullptr, nullptr, -1));          // raw-count cosine (recommender_similarity.cpp:99-104)
        } else {
            tok.clear(); val.clear();
            tok.reserve(it->second.size()); val.reserve(it->second.size());
            for (auto& kv : it->second) { tok.push_back(kv.first); val.push_back(kv.second); }
            PKCHECK(pk_set_idf(s.store, t, tok.data(), val.data(), (int64_t)tok.size()));
        }
        s.idf_fp[t] = fp;
        s.names[t] = cols[t];
    }
    return s;
}

// re-sends the adjacency rows a query on `user` reads: adj[user] and adj[f] for every friend f
// (recommender_graph.cpp:10-31,114-125; recommender_clubs.cpp:45-64).  The library keeps an override only
// where the live row differs from the uploaded one.
void refresh_adjacency(Shim& s, const Recommender& r, int user) {
    if (!r.adj_list) return;
    auto it = r.adj_list->find(user);
    if (it == r.adj_list->end()) { PKCHECK(pk_set_adj_row(s.store, user, nullptr, 0)); return; }
    PKCHECK(pk_set_adj_row(s.store, user, it->second.data(), (int64_t)it->second.size()));
    for (int f : it->second) {
        auto jt = r.adj_list->find(f);
        if (jt == r.adj_list->end()) PKCHECK(pk_set_adj_row(s.store, f, nullptr, 0));
        else PKCHECK(pk_set_adj_row(s.store, f, jt->second.data(), (int64_t)jt->second.size()));
    }
}

vector<pair<int, float>> run_mode(const Recommender& r, const vector<string>& cols, int mode, int user, int topk, int limit) {
    vector<pair<int, float>> out;
    if (!r.adj_list || !r.profiles) return out;                            // recommender_graph.cpp:36; clubs.cpp:13
    if (topk <= 0) return out;
    lock_guard<mutex> lk(g_mu);
    Shim& s = sync_state(r, cols);
    refresh_adjacency(s, r, user);
    vector<int32_t> id((size_t)topk);
    vector<float> sc((size_t)topk);
    int32_t cnt = 0, q = user;
    PKCHECK(pk_recommend(s.store, mode, 1, &q, topk, limit, id.data(), sc.data(), &cnt));
    out.reserve((size_t)cnt);
    for (int i = 0; i < cnt; ++i) out.emplace_back(id[i], sc[i]);
    return out;
}

}  // namespace

namespace {

// ---- ad-hoc pairs: two arbitrary UserProfile objects scored on the device ------------------------------
// profile_similarity and the private per-field helpers take profile OBJECTS, not map keys; they may not be
// members of `profiles` at all.  The pair is loaded as a 2-row world (uids 1 and 2) into a scratch store;
// the idf tables pushed are the caller's, restricted to the tokens the two profiles hold (a token absent
// from the caller's table stays absent -> idf 1.0f, recommender.cpp:78).
pk_store* g_scratch = nullptr;

struct AdhocOut { float fas = 0.f; vector<float> fields; };

// idf_of(t) returns the column's table or nullptr (no entry -> raw-count cosine)
template <class IdfOf>
AdhocOut adhoc_pair(const UserProfile& A, const UserProfile& B, int T, const Recommender* r, const vector<string>* cols, IdfOf idf_of) {
    if (!g_scratch) PKCHECK(pk_store_create(device_from_env(), &g_scratch));
    Flat f;
    f.begin();
    f.add_profile(A, 1, T);
    f.add_profile(B, 2, T);
    PKCHECK(pk_store_load_world(g_scratch, f.finish(T)));
    if (r && cols) push_normalizers(g_scratch, *r, *cols);
    vector<int32_t> tok;
    vector<float> val;
    for (int t = 0; t < T; ++t) {
        const unordered_map<int, float>* m = idf_of(t);
        if (!m) continue;
        tok.clear(); val.clear();
        for (int side = 0; side < 2; ++side) {
            const UserProfile& p = side ? B : A;
            if ((size_t)t >= p.token_cols.size()) continue;
            for (auto& kv : p.token_cols[t]) {
                auto it = m->find(kv.first);
                if (it == m->end()) continue;
                if (side == 1 && (size_t)t < A.token_cols.size() && A.token_cols[t].count(kv.first)) continue;   // already pushed
                tok.push_back(kv.first); val.push_back(it->second);
            }
        }
        PKCHECK(pk_set_idf(g_scratch, t, tok.data(), val.data(), (int64_t)tok.size()));
    }
    AdhocOut o;
    const int32_t a = 1, b = 2;
    PKCHECK(pk_pair_similarity(g_scratch, 1, &a, &b, &o.fas));
    o.fields.assign((size_t)3 + T, -1.f);
    PKCHECK(pk_pair_fields(g_scratch, 1, &a, &b, o.fields.data()));
    return o;
}

float field_or_zero(float v) { return v < 0.f ? 0.f : v; }   // the helpers return 0 for an absent field

}  // namespace

// ========================================================================================================
// Recommender members (reference include/recommender.h:17-71)
// ========================================================================================================

// src/recommender.cpp:9-16
Recommender::Recommender(const unordered_map<int, UserProfile>* profiles_in, const unordered_map<int, vector<int>>* al) {
    profiles = profiles_in;
    user_feats = nullptr;
    adj_list = al;
    total_users = profiles ? profiles->size() : 0;
    lock_guard<mutex> lk(g_mu);
    auto it = g_shims.find(this);              // the class has no destructor: an address can be reused
    if (it != g_shims.end()) { drop(it->second); g_shims.erase(it); }
}

// src/recommender.cpp:18-25.  The sparse-float `user_feats` branch has no live caller (only the dead
// evaluator.cpp:83) and no device path: every recommend call on such an instance returns empty.
Recommender::Recommender(const unordered_map<int, unordered_map<int, float>>* user_feats_in, const unordered_map<int, vector<int>>* al) {
    user_feats = user_feats_in;
    profiles = nullptr;
    adj_list = al;
    total_users = user_feats ? user_feats->size() : 0;
    lock_guard<mutex> lk(g_mu);
    auto it = g_shims.find(this);
    if (it != g_shims.end()) { drop(it->second); g_shims.erase(it); }
}

// src/recommender.cpp:27-41: plain member copies; the device is updated lazily by sync_state
void Recommender::set_field_normalizers(const unordered_map<string, pair<float, float>>& m) { field_normalizers = m; }
void Recommender::set_column_normalizers(const unordered_map<string, pair<float, float>>& m) { column_normalizers = m; }
void Recommender::set_text_columns(const vector<string>& cols) { text_columns_internal = cols; }
void Recommender::set_tfidf_index(const unordered_map<string, unordered_map<int, float>>& idf_map) { idf_per_col = idf_map; }

// src/recommender.cpp:43-66: df counted on the device, idf = logf(1 + N/(1+df)); the public host map is
// filled from the device table because callers copy it out (src/test.cpp:40)
void Recommender::compute_idf_from_profiles(const vector<string>& text_columns) {
    idf_per_col.clear();
    if (!profiles) return;
    total_users = profiles->size();
    lock_guard<mutex> lk(g_mu);
    Shim& s = g_shims[this];
    const int T = (int)text_columns.size();
    upload_world(s, *this, T);
    PKCHECK(pk_compute_idf(s.store, 0));
    vector<int32_t> tok;
    vector<float> val;
    for (int t = 0; t < T; ++t) {
        int64_t n = 0;
        PKCHECK(pk_idf_size(s.store, t, &n));
        tok.assign((size_t)max<int64_t>(n, 1), 0); val.assign((size_t)max<int64_t>(n, 1), 0.f);
        if (n > 0) PKCHECK(pk_get_idf(s.store, t, tok.data(), val.data(), nullptr, n));
        unordered_map<int, float> m;
        m.reserve((size_t)max<int64_t>(n, 0));
        for (int64_t i = 0; i < n; ++i) m.emplace(tok[i], val[i]);
        idf_per_col[text_columns[t]] = std::move(m);
    }
    // a later name in the list overwrites an earlier duplicate, as in the reference; fingerprints are taken
    // from the final host maps so that the next sync_state sees the device tables as current
    for (int t = 0; t < T; ++t) {
        s.idf_fp[t] = mix(idf_fingerprint(idf_per_col[text_columns[t]]), hash_str(text_columns[t]));
        s.names[t] = text_columns[t];
    }
}

// ---- recommend calls -------------------------------------------------------------------------------------
// src/recommender_graph.cpp:33-103
vector<pair<int, float>> Recommender::recommend_graph_registration(int user, int topk, int candidate_limit) const {
    return run_mode(*this, text_columns_internal, PK_MODE_GRAPH, user, topk, candidate_limit);
}
// src/recommender_graph.cpp:105-222
vector<pair<int, float>> Recommender::recommend_collaborative(int user, int topk, int candidate_limit) const {
    return run_mode(*this, text_columns_internal, PK_MODE_COLLAB, user, topk, candidate_limit);
}
// src/recommender_graph.cpp:224-227: an alias of recommend_graph_registration
vector<pair<int, float>> Recommender::recommend_by_interest(int user, int topk, int candidate_limit) const {
    return recommend_graph_registration(user, topk, candidate_limit);
}
// src/recommender_clubs.cpp:10-73 (candidate_limit is ignored there too)
vector<pair<int, float>> Recommender::recommend_clubs_collab(int user, int topk, int candidate_limit) const {
    return run_mode(*this, text_columns_internal, PK_MODE_CLUBS, user, topk, candidate_limit);
}
// src/recommender_graph.cpp:229-237
vector<pair<int, float>> Recommender::recommend_friends_graph(int user, int topk, int candidate_limit) const {
    return recommend_graph_registration(user, topk, candidate_limit);
}
vector<pair<int, float>> Recommender::recommend_friends_collab(int user, int topk, int candidate_limit) const {
    return recommend_collaborative(user, topk, candidate_limit);
}
vector<pair<int, float>> Recommender::recommend_friends_by_interest(int user, int topk, int candidate_limit) const {
    return recommend_by_interest(user, topk, candidate_limit);
}

// ---- batched calls (recommender_b200_batch.h) -------------------------------------------------------
// (the private column list is read through the friend struct of recommender_internal_sim.h)
namespace pokec_b200 {

int batch_mode_from_name(const string& name) {
    if (name == "graph" || name == "interest") return PK_MODE_GRAPH;      // recommender_graph.cpp:224-227: interest is an alias
    if (name == "collaborative") return PK_MODE_COLLAB;
    if (name == "clubs") return PK_MODE_CLUBS;
    if (name == "fullscan") return kBatchFullscan;
    return -1;
}

vector<vector<pair<int, float>>> recommend_batch(const Recommender& r, int pk_mode, const vector<int>& users, int topk, int candidate_limit) {
    vector<vector<pair<int, float>>> out(users.size());
    if (!r.adj_list || !r.profiles || topk <= 0 || users.empty()) return out;
    lock_guard<mutex> lk(g_mu);
    Shim& s = sync_state(r, RecommenderInternalSim::columns(r));
    for (int u : users) refresh_adjacency(s, r, u);
    const size_t nq = users.size();
    vector<int32_t> q(users.begin(), users.end()), id(nq * (size_t)topk), cnt(nq, 0);
    vector<float> sc(nq * (size_t)topk);
    if (pk_mode == kBatchFullscan)   // every profile is a candidate; the user and the existing friends are excluded (recommender_graph.cpp:46-52)
        PKCHECK(pk_fullscan_topk(s.store, (int)nq, q.data(), topk, PK_EXCLUDE_SELF_AND_ADJ, id.data(), sc.data(), cnt.data()));
    else
        PKCHECK(pk_recommend(s.store, pk_mode, (int)nq, q.data(), topk, candidate_limit, id.data(), sc.data(), cnt.data()));
    for (size_t i = 0; i < nq; ++i) {
        out[i].reserve((size_t)cnt[i]);
        for (int j = 0; j < cnt[i]; ++j) out[i].emplace_back(id[i * topk + j], sc[i * topk + j]);
    }
    return out;
}

// ---- the reference's containers <-> one world file (pk_world_save / pk_world_open) ----------------------
bool save_world_file(const Recommender& r, const string& path) {
    if (!r.profiles) return false;
    const vector<string>& cols = RecommenderInternalSim::columns(r);
    const int T = (int)cols.size();
    Flat f;
    flatten(r, T, f);
    normalizer_arrays(r, cols, f.nmean, f.nsd, f.nhas);
    pk_world w = *f.finish(T);
    w.norm_mean = f.nmean.data(); w.norm_sd = f.nsd.data(); w.norm_has = f.nhas.data();
    if (pk_world_save(&w, path.c_str()) != PK_OK) {
        fprintf(stderr, "[pokec_b200] save_world_file: %s\n", pk_last_error());
        return false;
    }
    return true;
}

int load_world_file(const string& path, unordered_map<int, UserProfile>& profiles, unordered_map<int, vector<int>>& adj_list,
                    vector<float>* norm_mean, vector<float>* norm_sd, vector<uint8_t>* norm_has) {
    pk_world_file* f = nullptr;
    pk_world w;
    if (pk_world_open(path.c_str(), &f, &w) != PK_OK) {
        fprintf(stderr, "[pokec_b200] load_world_file: %s\n", pk_last_error());
        return -1;
    }
    const int n = w.n_users, T = w.n_text_cols, m = w.n_adj_rows;
    profiles.clear();
    profiles.reserve((size_t)n);
    for (int i = 0; i < n; ++i) {
        UserProfile p;
        p.user_id = w.uid[i];
        p.public_flag = w.pub[i]; p.completion_percentage = w.comp[i]; p.gender = w.gender[i]; p.age = w.age[i];
        p.region_parts = {w.region[3 * i], w.region[3 * i + 1], w.region[3 * i + 2]};
        p.clubs.assign(w.clubs + w.clubs_ptr[i], w.clubs + w.clubs_ptr[i + 1]);
        p.friends.assign(w.friends + w.friends_ptr[i], w.friends + w.friends_ptr[i + 1]);
        p.token_cols.resize((size_t)T);
        for (int t = 0; t < T; ++t) {
            const int64_t b = w.tok_ptr[(int64_t)i * T + t], e = w.tok_ptr[(int64_t)i * T + t + 1];
            p.token_cols[t].reserve((size_t)(e - b));
            for (int64_t j = b; j < e; ++j) p.token_cols[t].emplace(w.tok_id[j], w.tok_tf[j]);
        }
        profiles.emplace(w.uid[i], std::move(p));
    }
    adj_list.clear();
    adj_list.reserve((size_t)m);
    for (int i = 0; i < m; ++i) adj_list.emplace(w.adj_uid[i], vector<int>(w.adj + w.adj_ptr[i], w.adj + w.adj_ptr[i + 1]));
    const int nf = PK_NUM_FIXED + T;
    if (norm_mean) norm_mean->assign(w.norm_mean, w.norm_mean + nf);
    if (norm_sd) norm_sd->assign(w.norm_sd, w.norm_sd + nf);
    if (norm_has) norm_has->assign(w.norm_has, w.norm_has + nf);
    pk_world_close(f);
    return T;
}

}  // namespace pokec_b200

// src/recommender_clubs.cpp:75-149.  Only caller is the dead evaluator.cpp:86:
// out of scope, no device path, returns no recommendations.
vector<pair<int, float>> Recommender::recommend_from_supernodes(int, const unordered_map<int, unordered_map<int, float>>&, int) const {
    static bool warned = false;
    if (!warned) { fprintf(stderr, "[pokec_b200] recommend_from_supernodes is not part of the accelerated path; returning no results\n"); warned = true; }
    return {};
}

// ---- pair similarity ---------------------------------------------------------------------------------------
// src/recommender_similarity.cpp:10-124
float Recommender::profile_similarity(const UserProfile& A, const UserProfile& B, const vector<string>& text_columns) const {
    lock_guard<mutex> lk(g_mu);
    // fast path: both objects are entries of `profiles` and the list is the one the store was built for
    if (profiles && &text_columns == &text_columns_internal) {
        auto ia = profiles->find(A.user_id), ib = profiles->find(B.user_id);
        if (ia != profiles->end() && ib != profiles->end() && &ia->second == &A && &ib->second == &B) {
            Shim& s = sync_state(*this, text_columns_internal);
            const int32_t a = ia->first, b = ib->first;
            float out = 0.f;
            PKCHECK(pk_pair_similarity(s.store, 1, &a, &b, &out));
            return out;
        }
    }
    const int T = (int)text_columns.size();
    if (T > 64) { fprintf(stderr, "[pokec_b200] fatal: %d text columns (at most 64 supported)\n", T); abort(); }
    AdhocOut o = adhoc_pair(A, B, T, this, &text_columns, [&](int t) -> const unordered_map<int, float>* {
        auto it = idf_per_col.find(text_columns[t]);                      // :99-104
        return it == idf_per_col.end() ? nullptr : &it->second;
    });
    return o.fas;
}

// src/recommender_similarity.cpp:126-128
float Recommender::profile_similarity(const UserProfile& A, const UserProfile& B) const {
    return profile_similarity(A, B, text_columns_internal);
}

// ---- private per-field helpers (declared in the header, so defined here; each is one device pair) ---------
// src/recommender.cpp:68-117
float Recommender::tfidf_cosine_for_column(const unordered_map<int, int>& A, const unordered_map<int, int>& B,
                                           const unordered_map<int, float>& idf_map) const {
    if (A.empty() || B.empty()) return 0.0f;
    UserProfile pa, pb;
    pa.token_cols.push_back(A);
    pb.token_cols.push_back(B);
    lock_guard<mutex> lk(g_mu);
    AdhocOut o = adhoc_pair(pa, pb, 1, nullptr, nullptr, [&](int) { return &idf_map; });
    return field_or_zero(o.fields[3]);
}

// src/recommender.cpp:119-128
float Recommender::vec_set_similarity(const vector<uint32_t>& A, const vector<uint32_t>& B) {
    if (A.empty() || B.empty()) return 0.0f;
    UserProfile pa, pb;
    pa.clubs = A;
    pb.clubs = B;
    lock_guard<mutex> lk(g_mu);
    AdhocOut o = adhoc_pair(pa, pb, 0, nullptr, nullptr, [](int) -> const unordered_map<int, float>* { return nullptr; });
    return field_or_zero(o.fields[1]);
}

// src/recommender.cpp:130-139
float Recommender::region_similarity_local(const array<int, 3>& A, const array<int, 3>& B) {
    UserProfile pa, pb;
    pa.region_parts = A;
    pb.region_parts = B;
    lock_guard<mutex> lk(g_mu);
    AdhocOut o = adhoc_pair(pa, pb, 0, nullptr, nullptr, [](int) -> const unordered_map<int, float>* { return nullptr; });
    return field_or_zero(o.fields[0]);
}

// src/recommender.cpp:141-163
float Recommender::cosine_counts_maps_local(const unordered_map<int, int>& A, const unordered_map<int, int>& B) {
    if (A.empty() || B.empty()) return 0.0f;
    UserProfile pa, pb;
    pa.token_cols.push_back(A);
    pb.token_cols.push_back(B);
    lock_guard<mutex> lk(g_mu);
    AdhocOut o = adhoc_pair(pa, pb, 1, nullptr, nullptr, [](int) -> const unordered_map<int, float>* { return nullptr; });
    return field_or_zero(o.fields[3]);
}
