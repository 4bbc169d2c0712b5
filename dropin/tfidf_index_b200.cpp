// tfidf_index_b200.cpp -- drop-in for the reference's src/tfidf_index.cpp (include/tfidf_index.h:10-21,
// header unchanged, compiled with -I <reference>/include).
//
// TFIDFIndex::build (src/tfidf_index.cpp:7-30) runs on the device through the C ABI: document frequencies
// are counted by the df kernel and idf = (float)log(1.0 + N/(1.0 + df)) is tabulated with the host libm
// (pk_compute_idf(store, double_log = 1)), so both the integer df table and the float idf table are
// bit-identical to the reference's.  weighted_cosine / compute_tfidf_vector are only reachable from the
// dead evaluator.cpp; they weight with the UNROUNDED double idf, which the
// device store does not keep, so they are defined (the header declares them) but not accelerated.
#include "tfidf_index.h"
#include "user_profile.h"

#include "pokec_b200.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace std;

namespace {
[[noreturn]] void die_tfidf(const char* what) {
    fprintf(stderr, "[pokec_b200] fatal: %s: %s\n", what, pk_last_error());
    abort();
}
#define PKCHECK_T(call) do { if ((call) != PK_OK) die_tfidf(#call); } while (0)
}  // namespace

void TFIDFIndex::build(const unordered_map<int, UserProfile>& profiles, const vector<string>& text_columns) {
    N = (int)profiles.size();
    const int T = (int)text_columns.size();
    doc_freqs.clear();
    doc_freqs.resize(text_columns.size());
    idf_per_col.clear();
    if (T > 64) { fprintf(stderr, "[pokec_b200] fatal: %d text columns (at most 64 supported)\n", T); abort(); }

    // flat token cells of every profile (only the text part of the world is needed for df)
    const int n = N;
    vector<int32_t> uid, zero((size_t)max(n, 1), 0), region((size_t)max(n, 1) * 3, -1), tok_id, tok_tf;
    vector<int64_t> ptr0((size_t)n + 1, 0), tok_ptr(1, 0);
    uid.reserve((size_t)n);
    for (auto& kv : profiles) {
        uid.push_back(kv.first);
        for (int t = 0; t < T; ++t) {
            if ((size_t)t < kv.second.token_cols.size())
                for (auto& pr : kv.second.token_cols[t]) { tok_id.push_back(pr.first); tok_tf.push_back(pr.second); }
            tok_ptr.push_back((int64_t)tok_id.size());
        }
    }
    pk_world w;
    memset(&w, 0, sizeof(w));
    w.n_users = n; w.n_text_cols = T;
    w.uid = uid.data(); w.pub = zero.data(); w.comp = zero.data(); w.gender = zero.data(); w.age = zero.data();
    w.region = region.data(); w.clubs_ptr = ptr0.data(); w.friends_ptr = ptr0.data();
    w.tok_ptr = tok_ptr.data(); w.tok_id = tok_id.data(); w.tok_tf = tok_tf.data();
    const char* e = getenv("POKEC_B200_DEVICE");
    pk_store* st = nullptr;
    PKCHECK_T(pk_store_create(e ? atoi(e) : 0, &st));
    PKCHECK_T(pk_store_load_world(st, &w));
    PKCHECK_T(pk_compute_idf(st, 1));
    vector<int32_t> tok, df;
    vector<float> val;
    for (int t = 0; t < T; ++t) {
        int64_t m = 0;
        PKCHECK_T(pk_idf_size(st, t, &m));
        tok.assign((size_t)max<int64_t>(m, 1), 0); df.assign(tok.size(), 0); val.assign(tok.size(), 0.f);
        if (m > 0) PKCHECK_T(pk_get_idf(st, t, tok.data(), val.data(), df.data(), m));
        unordered_map<int, float> idf;
        idf.reserve((size_t)max<int64_t>(m, 0));
        doc_freqs[t].reserve((size_t)max<int64_t>(m, 0));
        for (int64_t i = 0; i < m; ++i) { idf.emplace(tok[i], val[i]); doc_freqs[t].emplace(tok[i], df[i]); }
        idf_per_col[text_columns[t]] = std::move(idf);
    }
    pk_store_destroy(st);
}

float TFIDFIndex::weighted_cosine(const unordered_map<int, int>&, const unordered_map<int, int>&, int) const {
    static bool warned = false;
    if (!warned) { fprintf(stderr, "[pokec_b200] TFIDFIndex::weighted_cosine is not part of the accelerated path; returning 0\n"); warned = true; }
    return 0.0f;
}

void TFIDFIndex::compute_tfidf_vector(const UserProfile&, unordered_map<int, float>& out) const {
    static bool warned = false;
    if (!warned) { fprintf(stderr, "[pokec_b200] TFIDFIndex::compute_tfidf_vector is not part of the accelerated path; returning an empty vector\n"); warned = true; }
    out.clear();
}
