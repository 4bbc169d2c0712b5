// pk_host.cu -- host side of libpokec_b200.so: the C ABI of include/pokec_b200.h.
//
// Host work is limited to what the reference's Recommender also does outside the per-pair math:
// uid -> row lookup, walking the caller's adjacency to enumerate candidates in the reference's
// order (src/recommender_graph.cpp:10-31,114-125), planning query groups, and decoding results.
// Every similarity, every idf count and every top-k selection runs in the sm_100a kernels of
// pk_build.cuh / pk_scan.cuh.  There is no CPU scoring path.
#include "pokec_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstddef>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <atomic>
#include <thread>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "pk_layout.h"
#include "pk_build.cuh"
#include "pk_scan.cuh"
#include "pk_prefilter.cuh"
#include "pk_clubs.cuh"
#include "pk_pairs.cuh"
#include "pk_expand.cuh"

using namespace pk;

namespace {

thread_local std::string g_err = "";

int fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CK(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess)                                                                         \
            return fail(PK_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
    } while (0)

#define TRY(call)                 \
    do {                          \
        int rc_ = (call);         \
        if (rc_ != PK_OK) return rc_; \
    } while (0)

// grow-only device buffer
template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    int reserve(size_t n) {
        if (n <= cap) return PK_OK;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = n + n / 4 + 16;
        cudaError_t e = cudaMalloc((void**)&p, want * sizeof(T));
        if (e != cudaSuccess) return fail(PK_ERR_CUDA, "cudaMalloc(%zu bytes): %s", want * sizeof(T), cudaGetErrorString(e));
        cap = want;
        return PK_OK;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    ~DevBuf() { release(); }
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
};

int log2_table(uint64_t entries) {   // smallest L >= 1 with 2^L >= 2*entries
    int L = 1;
    while ((1ull << L) < 2 * entries) ++L;
    return L;
}

// host restatement of the reference's logistic / z (src/recommender_similarity.cpp:18-36); used ONLY to
// tabulate constants that depend on the normalizers, never on a pair of profiles.
double h_sigmoid(double x) {
    if (x >= 0) { double e = std::exp(-x); return 1.0 / (1.0 + e); }
    double e = std::exp(x);
    return e / (1.0 + e);
}

}  // namespace

struct FullscanJob {
    bool valid = false;
    int64_t gen = -1;                         // pk_store::gen the plan was built for
    int nq = 0, k = 0, K = 0, Kp = 0, exclude = 0, n_groups = 0;
    bool with_excl = false;
    size_t n_sm = 0, n_gl = 0;
    uint32_t max_smem_arena = 0;
    uint64_t arena_bytes = 0;
    int64_t pairs = 0;
    DevBuf<GroupDesc> groups;
    DevBuf<WorkItem> items;
    DevBuf<uint8_t> arena;
    DevBuf<int32_t> excl, q_slot0, q_nl, q_lane;
    DevBuf<uint64_t> keys_a, keys_b;
    // two-stage scan (pk_prefilter.cuh): pre-filter shortlists -> exact re-score -> fallback for unproven queries
    bool lean = false;
    int Ks = 0;
    size_t n_pre = 0;                         // pre-filter work items
    size_t n_rr_sm = 0, n_rr_gl = 0;          // exact re-score items (list mode over the shortlists)
    uint32_t max_smem_lean = 0;
    int64_t rr_pairs = 0;
    DevBuf<WorkItem> pre_items, rr_items;
    DevBuf<uint64_t> keys_pre, keys_short, keys_short_all, keys_rr;
    size_t n_rr = 0;
    // dynamic schedule of the pre-filter (PreParams): schedulable groups, {cursor, instances} per group
    DevBuf<int32_t> dyn_groups, dyn_state;
    int n_dyn = 0, n_slices = 1, sm_count = 148;
    int64_t cand_begin = 0, cand_end = 0;     // the shard the job was planned for
    DevBuf<int32_t> q_group, q_flag, q_flag0, group_flag, group_flag0, short_rows, rr_slot0, rr_nl, pre_slot0, pre_nl;
    // seed pass of the pruning thresholds: the rows of the shard with the most tokens (ascending row order)
    int n_seed = 0;
    uint32_t plan_fp = 0;                     // fingerprint of (queries, k, exclusion): part of the threshold tag
    DevBuf<int32_t> rs_a, rs_b, rs_id, rs_seg;  // exact re-score of the shortlists as (query, candidate) pairs: [nq][Ks]
    DevBuf<float> rs_f;
    DevBuf<int32_t> lean_fail;                // [group] the pre-filter's token table could not be built -> exact scan
};

struct pk_store {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int max_smem = 0;

    // ---- host-side view of the world -------------------------------------------------------
    int n = 0, T = 0;
    bool loaded = false;
    std::vector<int32_t> uid;
    std::vector<std::pair<int32_t, int32_t>> uid_sorted;   // (uid, row)
    std::vector<uint32_t> h_ncl, h_nfr, h_ntok;
    std::vector<int32_t> h_pub, h_gender, h_comp, h_age;   // scalar fields (pk_normalizers_from_pairs)
    std::vector<int32_t> adj_uid;
    std::vector<int64_t> adj_ptr;
    std::vector<int32_t> adj;
    std::unordered_map<int32_t, int32_t> adj_row;
    std::unordered_map<int32_t, std::vector<int32_t>> adj_override;

    // ---- device raw world ---------------------------------------------------------------------
    DevBuf<int32_t> d_uid, d_pub, d_comp, d_gender, d_age, d_region, d_tok_id, d_tok_tf;
    DevBuf<int64_t> d_clubs_ptr, d_friends_ptr, d_tok_ptr;
    DevBuf<uint32_t> d_clubs, d_friends;
    int64_t nnz = 0;

    // ---- idf ------------------------------------------------------------------------------------
    std::vector<uint8_t> col_state;   // 0 absent, 1 computed from profiles, 2 caller-supplied
    std::vector<std::vector<int32_t>> custom_tok;
    std::vector<std::vector<float>> custom_idf;
    int double_log = 0;
    bool idf_dirty = true;
    DevBuf<uint32_t> d_idf_key, d_idf_df;
    DevBuf<float> d_idf_val;
    int idf_log_cap = 0;
    DevBuf<uint8_t> d_col_has;

    // ---- normalizers ----------------------------------------------------------------------------
    std::vector<float> nmean, nsd;
    std::vector<uint8_t> nhas;
    bool norm_dirty = true;
    int max_comp = 0, max_age = 0;
    DevBuf<NormTables> d_norm;
    DevBuf<double> d_tab_comp, d_tab_age;

    // ---- packed store ---------------------------------------------------------------------------
    bool store_dirty = true;
    DevBuf<uint4> d_rec;
    DevBuf<uint64_t> d_rec_off;
    std::vector<uint64_t> h_rec_off;
    uint64_t store_u4 = 0;
    int64_t shard_begin = 0, shard_end = 0;
    // lean fp32 records of the pre-filter
    bool lean_dirty = true, tf_negative = false;
    bool wide_ids = false;   // a club / friend id >= 2^31: the pre-filter's combined set table cannot tag it
    DevBuf<uint4> d_lrec;
    DevBuf<uint64_t> d_lrec_off;
    DevBuf<uint2> d_lmask;
    uint64_t lean_u4 = 0;
    DevBuf<LeanNorm> d_lnorm;
    DevBuf<float> d_tab_comp32, d_tab_age32;
    float margin = 0.f;
    int64_t stat_fallback = 0, stat_prefilter_ns = 0, stat_rescore_ns = 0;
    bool rescore_time_pending = false;
    cudaEvent_t ev2 = nullptr, ev3 = nullptr;
    // phase marks of the two-stage scan (per-step breakdown of bench.py): start, after the table build, after the
    // shortlist merge | start of the second half, after gather-merge + proof, after the final merge
    cudaEvent_t pev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool phase_pending = false;
    int64_t stat_prep_ns = 0, stat_merge_ns = 0;

    // ---- per-call scratch (grow-only) -------------------------------------------------------------
    DevBuf<GroupDesc> d_groups;
    DevBuf<WorkItem> d_items;
    DevBuf<uint8_t> d_arena;
    DevBuf<int32_t> d_excl, d_cand_rows, d_i32a, d_i32b, d_i32c, d_i32d, d_i32e;
    DevBuf<int64_t> d_i64a;
    DevBuf<uint64_t> d_keys_a, d_keys_b, d_keys_c;
    DevBuf<float> d_f32a, d_f32b, d_fields;
    DevBuf<double> d_f64a;
    DevBuf<uint32_t> d_u32a;
    DevBuf<int> d_flag;
    DevBuf<ClubEdge> d_edges;
    // per-edge FAS cache: FAS(u, v) of every loaded adjacency entry u -> v, aligned with `adj`
    // (-1 where either side has no profile); valid until the normalizers, the idf tables or the world change
    DevBuf<float> d_edge_fas;
    std::vector<float> h_edge_fas;
    bool edge_cache_valid = false;
    // ---- device-resident graph for the batched list modes (pk_expand.cuh) -------------------------------------
    bool graph_dirty = true;
    int64_t max_deg = 0;
    DevBuf<int64_t> g_adj_ptr;
    DevBuf<int32_t> g_adj, g_adj_direct, g_row_direct, g_adj_suid, g_adj_srow, g_uid_s, g_uid_srow;
    DevBuf<unsigned long long> g_scratch;
    DevBuf<uint32_t> g_epoch;
    uint32_t g_tab_cap = 0;
    int g_warps = 0;
    DevBuf<int32_t> g_fr_slot, g_slot_row;           // [E] per adjacency entry of a queried user
    DevBuf<int32_t> g_qrow, g_cand, g_cnt, g_nslots, g_qoff, g_qdeg, g_cand_off, g_next;
    DevBuf<int64_t> g_pair_off, g_totals;
    DevBuf<uint32_t> g_ncl, g_cost, g_nfr, g_ntok;
    DevBuf<int32_t> g_group_off;
    DevBuf<unsigned long long> g_gtotals;
    DevBuf<int32_t> g_nedges, g_ntwo, g_edge_off, g_slot_begin;
    DevBuf<unsigned long long> g_ncon;
    FullscanJob job;
    // bumped by every call that changes what a prepared full-scan plan depends on (world, shard, normalizers, idf,
    // adjacency overrides): a plan stamped with an older value is refused by execute / finish / fetch / keys
    int64_t gen = 0;
    int64_t max_cell = 0;
    // pruning thresholds of the running full scan: one fixed allocation per store (its cudaIpc handle is handed to the
    // other ranks of a sharded scan, pk_fullscan_theta_export / _peers), zeroed per execution
    static constexpr size_t kThetaCap = (size_t)8192 * 32;
    DevBuf<unsigned long long> d_theta_x;
    const unsigned long long* peer_theta[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    void* peer_mapped[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // cudaIpcOpenMemHandle results
    int n_peers = 0;
    uint32_t exec_counter = 0;
    int reserve_sms = 0;              // SMs the persistent pre-filter leaves free (pk_store_set_option)
    DevBuf<int32_t> d_seed_rows;
    DevBuf<uint8_t> d_seed_skip;
    std::vector<int32_t> seed_cache;          // seed rows of the shard [seed_b, seed_e) with divisor seed_div
    int64_t seed_b = -1, seed_e = -1;
    int seed_div = 0;     // longest token cell of the loaded world (scales the pre-filter's error bound)
    bool scan_time_pending = false;
    int64_t stat_h2d = 0, stat_d2h = 0;
    int64_t last_idf_n = 0;

    // ---- stats ------------------------------------------------------------------------------------
    int64_t stat_launches = 0, stat_pairs = 0, stat_scan_ns = 0, stat_scan_launches = 0, stat_groups = 0;

    // direct uid -> row / adjacency-row maps (built when the uid range is at most a few times the user count):
    // candidate enumeration does one such lookup per neighbour, a binary search over 1.6 M uids costs ~1 us each
    int32_t uid_lo = 0;
    std::vector<int32_t> row_direct, adj_direct;
    int row_of(int32_t u) const {
        if (!row_direct.empty()) {
            const int64_t i = (int64_t)u - uid_lo;
            return (i >= 0 && i < (int64_t)row_direct.size()) ? row_direct[(size_t)i] : -1;
        }
        auto it = std::lower_bound(uid_sorted.begin(), uid_sorted.end(), std::make_pair(u, INT32_MIN));
        if (it == uid_sorted.end() || it->first != u) return -1;
        return it->second;
    }
    // the caller's adjacency row (with pk_set_adj_row overrides); returns false when there is no row
    bool adj_of(int32_t u, const int32_t** b, const int32_t** e) const {
        if (!adj_override.empty()) {
            auto io = adj_override.find(u);
            if (io != adj_override.end()) {
                *b = io->second.data();
                *e = io->second.data() + io->second.size();
                return true;
            }
        }
        if (!adj_direct.empty()) {
            const int64_t i = (int64_t)u - uid_lo;
            const int32_t r = (i >= 0 && i < (int64_t)adj_direct.size()) ? adj_direct[(size_t)i] : -1;
            if (r < 0) return false;
            *b = adj.data() + adj_ptr[r];
            *e = adj.data() + adj_ptr[r + 1];
            return true;
        }
        auto it = adj_row.find(u);
        if (it == adj_row.end()) return false;
        *b = adj.data() + adj_ptr[it->second];
        *e = adj.data() + adj_ptr[it->second + 1];
        return true;
    }
};

namespace {

template <class T>
int upload(pk_store* s, DevBuf<T>& buf, const T* src, size_t n) {
    TRY(buf.reserve(std::max<size_t>(n, 1)));
    if (n) CK(cudaMemcpyAsync(buf.p, src, n * sizeof(T), cudaMemcpyHostToDevice, s->stream));
    s->stat_h2d += (int64_t)(n * sizeof(T));
    return PK_OK;
}
template <class T>
int upload(pk_store* s, DevBuf<T>& buf, const std::vector<T>& v) {
    return upload(s, buf, v.data(), v.size());
}

RawWorld raw_world(pk_store* s) {
    RawWorld W;
    W.n = s->n; W.T = s->T;
    W.uid = s->d_uid.p; W.pub = s->d_pub.p; W.comp = s->d_comp.p; W.gender = s->d_gender.p; W.age = s->d_age.p;
    W.region = s->d_region.p;
    W.clubs_ptr = s->d_clubs_ptr.p; W.clubs = s->d_clubs.p;
    W.friends_ptr = s->d_friends_ptr.p; W.friends = s->d_friends.p;
    W.tok_ptr = s->d_tok_ptr.p; W.tok_id = s->d_tok_id.p; W.tok_tf = s->d_tok_tf.p;
    return W;
}

IdfTable idf_table(pk_store* s) {
    IdfTable I;
    I.key = s->d_idf_key.p; I.df = s->d_idf_df.p; I.idf = s->d_idf_val.p; I.log_cap = s->idf_log_cap;
    return I;
}

// ---- idf table ------------------------------------------------------------------------------------
int rebuild_idf(pk_store* s) {
    if (!s->idf_dirty) return PK_OK;
    const int T = s->T;
    bool any_computed = false;
    uint64_t custom_entries = 0;
    std::vector<uint8_t> computed(std::max(T, 1), 0), has(std::max(T, 1), 0);
    for (int t = 0; t < T; ++t) {
        computed[t] = s->col_state[t] == 1;
        has[t] = s->col_state[t] != 0;
        any_computed |= computed[t] != 0;
        if (s->col_state[t] == 2) custom_entries += s->custom_tok[t].size();
    }
    TRY(upload(s, s->d_col_has, has));
    uint64_t want = custom_entries + 1024;
    if (any_computed) want += (uint64_t)s->nnz / 2;
    int log_cap = std::max(10, log2_table(want) - 1);
    DevBuf<uint8_t> d_computed;
    TRY(upload(s, d_computed, computed));
    TRY(s->d_flag.reserve(4));
    for (;; ++log_cap) {
        if (log_cap > 31) return fail(PK_ERR_UNSUPPORTED, "idf table would exceed 2^31 slots");
        const size_t cap = (size_t)1 << log_cap;
        TRY(s->d_idf_key.reserve(cap));
        TRY(s->d_idf_df.reserve(cap));
        TRY(s->d_idf_val.reserve(cap));
        s->idf_log_cap = log_cap;
        CK(cudaMemsetAsync(s->d_idf_key.p, 0xFF, cap * 4, s->stream));
        CK(cudaMemsetAsync(s->d_idf_df.p, 0, cap * 4, s->stream));
        CK(cudaMemsetAsync(s->d_idf_val.p, 0, cap * 4, s->stream));
        CK(cudaMemsetAsync(s->d_flag.p, 0, sizeof(int), s->stream));
        IdfTable I = idf_table(s);
        if (any_computed && s->n > 0) {
            const int64_t cells = (int64_t)s->n * T;
            df_count_kernel<<<(unsigned)((cells + 255) / 256), 256, 0, s->stream>>>(raw_world(s), I, d_computed.p, s->d_flag.p);
            ++s->stat_launches;
            // idf(df) evaluated with the host libm: bit-identical to src/recommender.cpp:61 (logf, float)
            // or src/tfidf_index.cpp:25 (double log, rounded to float)
            std::vector<float> idf_of_df((size_t)s->n + 1);
            const size_t total_users = (size_t)s->n;
            for (size_t d = 0; d <= total_users; ++d) {
                if (s->double_log) idf_of_df[d] = (float)std::log(1.0 + (double)(int)total_users / (1.0 + (double)(int)d));
                else idf_of_df[d] = logf(1.0f + (float)total_users / (1.0f + (float)(int)d));
            }
            TRY(upload(s, s->d_f32a, idf_of_df));
            CK(cudaStreamSynchronize(s->stream));   // idf_of_df is a local
            idf_fill_kernel<<<(unsigned)((cap + 255) / 256), 256, 0, s->stream>>>(I, s->d_f32a.p, (int64_t)s->n);
            ++s->stat_launches;
        }
        for (int t = 0; t < T; ++t) {
            if (s->col_state[t] != 2 || s->custom_tok[t].empty()) continue;
            const int64_t m = (int64_t)s->custom_tok[t].size();
            TRY(upload(s, s->d_i32a, s->custom_tok[t]));
            TRY(upload(s, s->d_f32a, s->custom_idf[t]));
            idf_custom_kernel<<<(unsigned)((m + 255) / 256), 256, 0, s->stream>>>(I, t, s->d_i32a.p, s->d_f32a.p, m, s->d_flag.p);
            ++s->stat_launches;
            CK(cudaStreamSynchronize(s->stream));
        }
        int overflow = 0;
        CK(cudaMemcpyAsync(&overflow, s->d_flag.p, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        if (!overflow) break;
    }
    s->idf_dirty = false;
    s->store_dirty = true;
    return PK_OK;
}

// ---- normalizer constants ------------------------------------------------------------------------
int rebuild_norm(pk_store* s) {
    if (!s->norm_dirty) return PK_OK;
    const int T = s->T, nf = kNumFixed + T;
    NormTables N;
    memset(&N, 0, sizeof(N));
    auto z = [&](int f, double sv) -> double {
        if (N.use_norm[f]) return (sv - N.mean[f]) / N.sd[f];
        return 6.0 * (sv - 0.5);
    };
    for (int f = 0; f < nf; ++f) {
        N.mean[f] = (double)s->nmean[f];
        N.sd[f] = (double)s->nsd[f];
        N.use_norm[f] = (s->nhas[f] && (double)s->nsd[f] > 0.0) ? 1 : 0;
        N.sig0[f] = h_sigmoid(z(f, 0.0));
        N.sig1[f] = h_sigmoid(z(f, 1.0));
    }
    N.total_possible = nf;
    for (int u = 0; u <= nf; ++u) N.ftab[u] = (double)u / (double)nf;
    for (int a = 0; a < 4; ++a)
        for (int b = 0; b < 4; ++b)
            for (int m = 0; m < 4; ++m) {
                double v = 0.0;
                if (a > 0 && b > 0) {
                    float sv = (float)((double)m / (std::sqrt((double)a) * std::sqrt((double)b)));
                    v = h_sigmoid(z(4, (double)sv));
                }
                N.region[(a * 4 + b) * 4 + m] = v;
            }
    N.dim_comp = std::min(std::max(s->max_comp, 0), 255) + 1;
    N.dim_age = std::min(std::max(s->max_age, 0), 255) + 1;
    for (int which = 0; which < 2; ++which) {
        const int dim = which ? N.dim_age : N.dim_comp;
        const int f = which ? 3 : 2;
        std::vector<double> tab((size_t)dim * dim, 0.0);
        for (int c = 1; c < dim; ++c)
            for (int q = 1; q < dim; ++q) {
                const int amin = std::min(c, q), amax = std::max(c, q);
                tab[(size_t)c * dim + q] = h_sigmoid(z(f, (double)amin / (double)amax));
            }
        TRY(upload(s, which ? s->d_tab_age : s->d_tab_comp, tab));
        std::vector<float> tab32(tab.begin(), tab.end());
        TRY(upload(s, which ? s->d_tab_age32 : s->d_tab_comp32, tab32));
        CK(cudaStreamSynchronize(s->stream));
    }
    {   // fp32 constants of the pre-filter and the bound on its error (pk_prefilter.cuh)
        std::vector<LeanNorm> Lv(1);
        LeanNorm& L = Lv[0];
        memset(&L, 0, sizeof(L));
        const double log2e = 1.4426950408889634;
        double worst = 1.5;                 // default z = 6(s - 0.5): |d sigma / d ln s| <= 0.25 * 6
        for (int f = 0; f < nf; ++f) {
            double a, b;
            if (N.use_norm[f]) {
                a = -log2e / N.sd[f]; b = log2e * N.mean[f] / N.sd[f];
                worst = std::max(worst, 0.25 + 0.25 * std::fabs(N.mean[f]) / N.sd[f]);
            } else { a = -6.0 * log2e; b = 3.0 * log2e; }
            L.fld[f] = make_float4((float)a, (float)b, (float)N.sig0[f], (float)N.sig1[f]);
        }
        for (int i = 0; i < 64; ++i) L.region[i] = (float)N.region[i];
        for (int j = 0; j < kMaxT / 8; ++j)
            for (int bits = 0; bits < 256; ++bits) {
                double acc = 0.0;
                for (int i = 0; i < 8; ++i)
                    if (((bits >> i) & 1) && 8 * j + i < T) acc += N.sig0[kNumFixed + 8 * j + i];
                L.lut[j][bits] = (float)acc;
            }
        L.inv_total = (float)(1.0 / (double)nf);
        // Error bound of the fp32 score.  A text cosine is a sum of m <= min(nnz_q, nnz_c) <= max_cell positive
        // products of two normalised weights: each weight is one rounding of an fp64 quotient (2^-24 relative), each
        // fmaf adds at most 2^-24 of the partial sum, so the cosine is off by at most (m + 2) 2^-24 relative --
        // 4.2e-6 for the 64 products the bound was first written for, scaled here by the LONGEST cell of the loaded
        // world (a 300-token cell gives 1.8e-5).  The set similarities (three roundings) are far inside that.
        // sigma(z(s)) moves by at most `worst` per unit of relative error of s; ex2.approx / rcp.approx, the fp32 sum
        // of <= 56 sigmas and the two approximate divisions add < 4e-6 absolute (5e-6 taken); the mean over fields
        // cannot exceed the worst field; dFAS/dS <= 2; safety factor 2.  lean_usable() gives the pre-filter up
        // (exact scan on every pair) when the bound exceeds 1e-3.
        const double rel_s = std::max(4.2e-6, ((double)s->max_cell + 6.0) * 5.9604645e-8);
        L.margin = (float)(2.0 * 2.0 * (worst * rel_s + 5e-6));
        L.dim_comp = N.dim_comp; L.dim_age = N.dim_age;
        {   // sigma is monotone in s on [0, 1], so a field's largest contribution is at s = 0 or s = 1
            auto up = [](double v) { return (float)(v * (1.0 + 4e-7) + 1e-7); };
            double mt = 0.0;
            for (int t = 0; t < T; ++t) mt = std::max(mt, std::max(N.sig0[kNumFixed + t], N.sig1[kNumFixed + t]));
            L.ub_text = up(mt);
            L.ub_clubs = up(std::max(N.sig0[5], N.sig1[5]));
            L.ub_friends = up(std::max(N.sig0[6], N.sig1[6]));
        }
        s->margin = L.margin;
        TRY(s->d_lnorm.reserve(1));
        CK(cudaMemcpyAsync(s->d_lnorm.p, &L, sizeof(L), cudaMemcpyHostToDevice, s->stream));
        CK(cudaStreamSynchronize(s->stream));
    }
    TRY(s->d_norm.reserve(1));
    CK(cudaMemcpyAsync(s->d_norm.p, &N, sizeof(N), cudaMemcpyHostToDevice, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    s->norm_dirty = false;
    return PK_OK;
}

// ---- packed record store ---------------------------------------------------------------------------
int rebuild_store(pk_store* s) {
    TRY(rebuild_idf(s));
    if (!s->store_dirty) return PK_OK;
    const int n = s->n;
    s->h_rec_off.assign((size_t)n + 1, 0);
    if (n > 0) {
        TRY(s->d_u32a.reserve((size_t)2 * n));
        const unsigned blocks = (unsigned)(((int64_t)n * 32 + 255) / 256);
        record_size_kernel<<<blocks, 256, 0, s->stream>>>(raw_world(s), s->d_u32a.p, s->d_u32a.p + n);
        ++s->stat_launches;
        std::vector<uint32_t> sz((size_t)2 * n);
        CK(cudaMemcpyAsync(sz.data(), s->d_u32a.p, sz.size() * 4, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        s->h_ntok.assign(sz.begin() + n, sz.end());
        for (int r = 0; r < n; ++r) s->h_rec_off[r + 1] = s->h_rec_off[r] + sz[r];
    }
    s->store_u4 = s->h_rec_off[n];
    TRY(s->d_rec.reserve(std::max<uint64_t>(s->store_u4, 1)));
    TRY(upload(s, s->d_rec_off, s->h_rec_off));
    if (n > 0) {
        const unsigned blocks = (unsigned)(((int64_t)n * 32 + 255) / 256);
        pack_kernel<<<blocks, 256, 0, s->stream>>>(raw_world(s), idf_table(s), s->d_col_has.p, s->d_rec_off.p, s->d_rec.p);
        ++s->stat_launches;
    }
    CK(cudaStreamSynchronize(s->stream));
    CK(cudaGetLastError());
    s->store_dirty = false;
    s->lean_dirty = true;
    return PK_OK;
}

// lean fp32 records of the pre-filter, derived from the packed store
int rebuild_lean(pk_store* s) {
    if (!s->lean_dirty) return PK_OK;
    const int n = s->n;
    std::vector<uint64_t> off((size_t)n + 1, 0);
    for (int r = 0; r < n; ++r) off[r + 1] = off[r] + lean_record_u4(s->h_ncl[r], s->h_nfr[r], s->h_ntok[r]);
    s->lean_u4 = off[n];
    TRY(s->d_lrec.reserve(s->lean_u4 + 16));      // + 256 B: the pre-filter reads the first 32 set words of a record speculatively
    TRY(upload(s, s->d_lrec_off, off));
    TRY(s->d_lmask.reserve((size_t)std::max(n, 1)));
    if (n > 0) {
        const unsigned blocks = (unsigned)(((int64_t)n * 32 + 255) / 256);
        lean_pack_kernel<<<blocks, 256, 0, s->stream>>>(n, s->d_rec.p, s->d_rec_off.p, s->d_lrec_off.p, s->d_lrec.p, s->d_lmask.p);
        ++s->stat_launches;
    }
    CK(cudaStreamSynchronize(s->stream));
    CK(cudaGetLastError());
    s->lean_dirty = false;
    return PK_OK;
}

int ensure_ready(pk_store* s) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    CK(cudaSetDevice(s->device));
    TRY(rebuild_store(s));
    TRY(rebuild_norm(s));
    return PK_OK;
}

uint32_t align16(uint32_t x) { return (x + 15u) & ~15u; }

// ---- group planning --------------------------------------------------------------------------------

struct Plan {
    std::vector<GroupDesc> groups;
    std::vector<int32_t> excl;    // concatenated exclusion ids
    uint64_t arena_bytes = 0;
    uint32_t max_smem_arena = 0;  // over groups that fit shared memory
    uint32_t max_smem_lean = 0;   // same for the pre-filter's tables
};

// lays out the tables of a group whose members are rows[0..n_lanes); excl_n = exclusion ids per lane.
// with_lean adds the fp32 tables of the pre-filter in front of the block (pk_layout.h, GroupDesc).
void layout_group(pk_store* s, GroupDesc& g, const uint32_t* excl_n, bool with_lean = false) {
    uint64_t ncl = 0, nfr = 0, ntok = 0, nex = 0;
    for (int l = 0; l < g.n_lanes; ++l) {
        const int r = g.row[l];
        if (r < 0) continue;
        ncl += s->h_ncl[r]; nfr += s->h_nfr[r]; ntok += s->h_ntok[r];
        if (excl_n) nex += excl_n[l];
    }
    g.log_sc = log2_table(ncl); g.log_sf = log2_table(nfr); g.log_sa = log2_table(nex); g.log_st = log2_table(ntok);
    uint32_t off = 0;
    g.off_text2 = g.off_base = g.off_vals32 = g.off_sets2 = 0; g.log_cb = g.log_cs = 0;
    if (with_lean) {
        g.log_cb = std::max(1, g.log_st - 1);
        g.off_text2 = off; off += (16u << g.log_cb);
        g.off_base = off; off += align16((uint32_t)ntok * 8u);
        g.off_vals32 = off; off += align16((uint32_t)ntok * 4u);
        g.log_cs = std::max(1, log2_table(ncl + nfr) - 1);
        g.off_sets2 = off; off += (16u << g.log_cs);
    }
    g.exact_begin = off;
    g.off_adj = off; off += (8u << g.log_sa);
    off = align16(off);
    g.lean_end = with_lean ? off : 0u;
    g.off_clubs = off; off += (8u << g.log_sc);
    g.off_friends = off; off += (8u << g.log_sf);
    off = align16(off);
    g.off_nq = off; off += align16((uint32_t)s->T * 32u * 8u);
    g.off_text = off; off += (16u << g.log_st);
    g.off_vals = off; off += align16((uint32_t)ntok * 8u);
    g.n_vals = (uint32_t)ntok;
    g.arena_bytes = align16(off);
}

uint32_t scan_fixed_smem(int Kp) {
    const uint32_t norm = (uint32_t)((sizeof(NormTables) + 15) & ~15u);
    return norm + 32u * Kp * 8u + 32u * 8u + 32u * 4u + 16u;
}

int upload_plan(pk_store* s, Plan& plan, bool with_excl) {
    uint64_t off = 0;
    for (auto& g : plan.groups) {
        g.arena_off = (uint32_t)off;
        off += g.arena_bytes;
        if (off > 0xF0000000ull) return fail(PK_ERR_UNSUPPORTED, "query batch too large for one arena (split the batch)");
    }
    plan.arena_bytes = off;
    TRY(s->d_arena.reserve(std::max<uint64_t>(off, 16)));
    CK(cudaMemsetAsync(s->d_arena.p, 0, std::max<uint64_t>(off, 16), s->stream));
    TRY(upload(s, s->d_groups, plan.groups));
    if (with_excl) TRY(upload(s, s->d_excl, plan.excl));
    if (!plan.groups.empty()) {
        PrepParams P;
        P.rec = s->d_rec.p; P.rec_off = s->d_rec_off.p; P.groups = s->d_groups.p; P.arena = s->d_arena.p;
        P.excl_ids = with_excl ? s->d_excl.p : nullptr;
        P.lean_fail = nullptr;
        group_prep_kernel<<<(unsigned)plan.groups.size(), kPrepThreads, 0, s->stream>>>(P);
        ++s->stat_launches;
        CK(cudaGetLastError());
    }
    return PK_OK;
}

template <int EPI, bool FIELDS>
int launch_scan_t(pk_store* s, ScanParams P, const WorkItem* d_items, size_t n_items, bool sm, uint32_t smem_bytes) {
    if (n_items == 0) return PK_OK;
    P.items = d_items;
    if (sm) {
        auto kern = fas_scan_kernel<true, EPI, FIELDS>;
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
        kern<<<(unsigned)n_items, kScanThreads, smem_bytes, s->stream>>>(P);
    } else {
        auto kern = fas_scan_kernel<false, EPI, FIELDS>;
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
        kern<<<(unsigned)n_items, kScanThreads, smem_bytes, s->stream>>>(P);
    }
    CK(cudaGetLastError());
    ++s->stat_launches;
    return PK_OK;
}

// launches the scan over device-resident items: first the n_sm items whose group tables are staged in
// shared memory, then the n_gl items whose tables stay in global memory.  Asynchronous; the scan time is
// read from the events by collect_scan_time() after the caller synchronised.
template <int EPI, bool FIELDS>
int launch_scan(pk_store* s, ScanParams P, const WorkItem* d_items, size_t n_sm, size_t n_gl, uint32_t max_smem_arena) {
    const uint32_t fixed = scan_fixed_smem(EPI == EPI_TOPK ? P.Kp : 0);
    CK(cudaEventRecord(s->ev0, s->stream));
    P.smem_arena_bytes = max_smem_arena;
    TRY((launch_scan_t<EPI, FIELDS>(s, P, d_items, n_sm, true, max_smem_arena + fixed)));
    P.smem_arena_bytes = 0;
    TRY((launch_scan_t<EPI, FIELDS>(s, P, d_items + n_sm, n_gl, false, fixed)));
    CK(cudaEventRecord(s->ev1, s->stream));
    s->scan_time_pending = true;
    s->stat_scan_launches += (n_sm > 0) + (n_gl > 0);
    return PK_OK;
}

// the same two launches without the scan-time events (re-score / fallback passes of the two-stage scan)
template <int EPI, bool FIELDS>
int launch_scan_plain(pk_store* s, ScanParams P, const WorkItem* d_items, size_t n_sm, size_t n_gl, uint32_t max_smem_arena) {
    const uint32_t fixed = scan_fixed_smem(EPI == EPI_TOPK ? P.Kp : 0);
    P.smem_arena_bytes = max_smem_arena;
    TRY((launch_scan_t<EPI, FIELDS>(s, P, d_items, n_sm, true, max_smem_arena + fixed)));
    P.smem_arena_bytes = 0;
    TRY((launch_scan_t<EPI, FIELDS>(s, P, d_items + n_sm, n_gl, false, fixed)));
    return PK_OK;
}

int collect_scan_time(pk_store* s) {
    if (!s->scan_time_pending) return PK_OK;
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
    s->stat_scan_ns += (int64_t)((double)ms * 1e6);
    s->scan_time_pending = false;
    if (s->rescore_time_pending) {
        CK(cudaEventElapsedTime(&ms, s->ev2, s->ev3));
        s->stat_rescore_ns = (int64_t)((double)ms * 1e6);
        s->rescore_time_pending = false;
        if (s->phase_pending) {
            float a = 0.f, b = 0.f, c = 0.f, d = 0.f;
            CK(cudaEventElapsedTime(&a, s->pev[0], s->pev[1]));      // arena clear + group_prep_kernel
            CK(cudaEventElapsedTime(&b, s->ev1, s->pev[2]));         // merge of the instances' shortlists
            CK(cudaEventElapsedTime(&c, s->pev[3], s->ev2));         // merge of the gathered shortlists + proof
            CK(cudaEventElapsedTime(&d, s->ev3, s->pev[5]));         // final merge
            s->stat_prep_ns = (int64_t)((double)a * 1e6);
            s->stat_merge_ns = (int64_t)(((double)b + c + d) * 1e6);
            s->phase_pending = false;
        }
    }
    return PK_OK;
}

// orders the items (shared-memory groups first) and uploads them into `buf`
int upload_items(pk_store* s, const Plan& plan, const std::vector<WorkItem>& items, DevBuf<WorkItem>& buf, size_t* n_sm,
                 size_t* n_gl) {
    std::vector<WorkItem> all;
    all.reserve(items.size());
    for (const auto& it : items) if (plan.groups[it.group].in_smem) all.push_back(it);
    *n_sm = all.size();
    for (const auto& it : items) if (!plan.groups[it.group].in_smem) all.push_back(it);
    *n_gl = all.size() - *n_sm;
    return upload(s, buf, all);
}

// upload + launch + synchronise (the list / pair paths)
template <int EPI, bool FIELDS>
int run_scan(pk_store* s, const Plan& plan, ScanParams P, const std::vector<WorkItem>& items) {
    size_t n_sm = 0, n_gl = 0;
    TRY(upload_items(s, plan, items, s->d_items, &n_sm, &n_gl));
    TRY((launch_scan<EPI, FIELDS>(s, P, s->d_items.p, n_sm, n_gl, plan.max_smem_arena)));
    CK(cudaStreamSynchronize(s->stream));
    return collect_scan_time(s);
}

ScanParams base_scan_params(pk_store* s, int K) {
    ScanParams P;
    memset(&P, 0, sizeof(P));
    P.rec = s->d_rec.p; P.rec_off = s->d_rec_off.p; P.groups = s->d_groups.p; P.arena = s->d_arena.p;
    P.norm = s->d_norm.p; P.tab_comp = s->d_tab_comp.p; P.tab_age = s->d_tab_age.p;
    P.T = s->T; P.K = K; P.Kp = (K + 31) / 32 * 32; P.exclude_mode = PK_EXCLUDE_NONE;
    P.nf = 3 + s->T + 2;
    P.group_lo = 0; P.group_hi = 0x7fffffff;
    P.want_smem = -1;
    return P;
}

uint32_t lean_fixed_smem(int Ks) {
    const uint32_t norm = (uint32_t)((sizeof(LeanNorm) + 15) & ~15u);
    // lists, tau, locks, theta, level-1 tables {l1q 32 x 16, l1f 32 x 32, l1b 32 x 4}, next / pick, per-warp hit staging
    return norm + 32u * Ks * 8u + 32u * 8u + 32u * 4u + 32u * 4u + (512u + 1024u + 128u + 896u) + 16u + (uint32_t)kPreWarps * 32u * 16u;
}

void finish_group(pk_store* s, Plan& plan, GroupDesc& g, const uint32_t* excl_n, int Kp, bool with_lean = false, int Ks = 0) {
    layout_group(s, g, excl_n, with_lean);
    const uint32_t budget = (uint32_t)s->max_smem - scan_fixed_smem(Kp) - 256u;
    g.in_smem = (g.arena_bytes - g.exact_begin) <= budget;
    if (g.in_smem) plan.max_smem_arena = std::max(plan.max_smem_arena, g.arena_bytes - g.exact_begin);
    g.lean_in_smem = 0;
    if (with_lean) {
        const uint32_t lbudget = (uint32_t)s->max_smem - lean_fixed_smem(Ks) - 256u;
        g.lean_in_smem = g.lean_end <= lbudget;
        if (g.lean_in_smem) plan.max_smem_lean = std::max(plan.max_smem_lean, g.lean_end);
    }
    plan.groups.push_back(g);
}

// merge_topk_kernel launch: sort buffer sized for K (power of two >= 2K, at least 2048 keys)
int launch_merge(pk_store* s, MergeParams M, int n_queries) {
    if (n_queries <= 0) return PK_OK;
    int cap = 2048;
    while (cap < 2 * M.K) cap <<= 1;
    M.cap = cap;
    const size_t smem = ((size_t)cap + (size_t)M.K) * 8;
    if (smem > (size_t)s->max_smem) return fail(PK_ERR_UNSUPPORTED, "top-k of %d exceeds the selection kernel's shared memory", M.K);
    if (smem > 48 * 1024) CK(cudaFuncSetAttribute(merge_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    merge_topk_kernel<<<(unsigned)n_queries, kMergeThreads, smem, s->stream>>>(M);
    ++s->stat_launches;
    CK(cudaGetLastError());
    return PK_OK;
}

// top-k of a flat key slice per user (collaborative / clubs / wide graph epilogue): keys[begin[u] .. begin[u+1])
int launch_slice_topk(pk_store* s, const uint64_t* keys, const int32_t* begin, int nu, int K, uint64_t* out_keys,
                      const int32_t* cnt = nullptr) {
    MergeParams M;
    memset(&M, 0, sizeof(M));
    M.keys = keys; M.seg_begin = begin; M.seg_cnt = cnt; M.nq = nu; M.K = K; M.out_keys = out_keys;
    return launch_merge(s, M, nu);
}

void decode_keys(const uint64_t* keys, int nq, int k, int32_t* out_id, float* out_score, int32_t* out_cnt) {
    for (int q = 0; q < nq; ++q) {
        int c = 0;
        for (int i = 0; i < k; ++i) {
            const uint64_t key = keys[(size_t)q * k + i];
            if (key == 0ull) { out_id[(size_t)q * k + i] = -1; out_score[(size_t)q * k + i] = 0.0f; continue; }
            out_id[(size_t)q * k + i] = key_id(key);
            uint32_t b = key_score_bits(key);
            float f;
            memcpy(&f, &b, 4);
            out_score[(size_t)q * k + i] = f;
            ++c;
        }
        out_cnt[q] = c;
    }
}

int check_topk_args(pk_store* s, int nq, const int32_t* q_uid, int k, int max_k = kMaxK) {
    if (nq < 0 || k < 0 || (nq > 0 && !q_uid)) return fail(PK_ERR_ARG, "bad nq/k/q_uid");
    if (k > max_k) return fail(PK_ERR_UNSUPPORTED, "k = %d exceeds the supported maximum %d", k, max_k);
    (void)s;
    return PK_OK;
}

template <bool FIELDS>
int pairs_launch(pk_store* s, int64_t n, const int32_t* d_a_row, const int32_t* d_b_row, float* d_out_f, float* d_out_fields,
                 const int32_t* d_key_id, uint64_t* d_out_keys, bool timed = true);

// ---- full scan ---------------------------------------------------------------------------------------
// prepare: host planning + upload (queries -> groups -> work items); execute: arena clear, group_prep,
// scan, merge -- all asynchronous on the store's stream, device-resident inputs only; fetch: D2H + decode.
int job_ok(pk_store* s) {
    if (!s->job.valid) return fail(PK_ERR_STATE, "pk_fullscan_prepare has not been called");
    if (s->job.gen != s->gen)
        return fail(PK_ERR_STATE, "the prepared full scan is stale: the world, shard, normalizers, idf or adjacency changed after "
                                  "pk_fullscan_prepare (prepare again)");
    return PK_OK;
}

bool lean_usable(pk_store* s, int K, int* Ks_out) {
    const char* e = getenv("POKEC_B200_NO_PREFILTER");
    if (e && e[0] == '1') return false;
    if (s->tf_negative || s->wide_ids) return false;
    if (!(s->margin < 1e-3f)) return false;     // cells so long that the fp32 bound is useless: exact scan on every pair
    for (int t = 0; t < s->T; ++t)
        if (s->col_state[t] == 2)
            for (float v : s->custom_idf[t]) if (!(v >= 0.0f)) return false;   // negative weights break the error bound
    int Ks = (2 * K + 28 + 31) / 32 * 32;
    if (Ks > kMaxK) Ks = kMaxK;
    if (Ks < K + 16) return false;
    *Ks_out = Ks;
    return true;
}

int fullscan_prepare(pk_store* s, int nq, const int32_t* q_uid, int k, int exclude_mode) {
    FullscanJob& J = s->job;
    J.valid = false;
    J.nq = nq; J.k = k; J.exclude = exclude_mode;
    {
        uint32_t h = 2166136261u;
        auto mixin = [&](uint32_t v) { h = (h ^ v) * 16777619u; };
        mixin((uint32_t)nq); mixin((uint32_t)k); mixin((uint32_t)exclude_mode);
        for (int q = 0; q < nq; ++q) mixin((uint32_t)q_uid[q]);
        J.plan_fp = h;
    }
    const int K = std::max(k, 1), Kp = (K + 31) / 32 * 32;
    J.K = K; J.Kp = Kp;
    J.n_groups = 0; J.n_sm = J.n_gl = 0; J.pairs = 0;
    J.lean = false; J.Ks = 0; J.n_pre = 0; J.n_rr_sm = J.n_rr_gl = 0; J.rr_pairs = 0;
    TRY(J.keys_b.reserve((size_t)std::max(nq, 1) * K));
    const int64_t n_cand = s->shard_end - s->shard_begin;
    if (nq == 0 || k == 0 || n_cand <= 0) { J.valid = true; J.gen = s->gen; return PK_OK; }
    int Ks = 0;
    const bool lean = lean_usable(s, K, &Ks);
    if (lean) TRY(rebuild_lean(s));

    // group the queries that have a profile, in caller order
    Plan plan;
    std::vector<int32_t> q_slot0(nq, 0), q_nl(nq, 0), q_lane(nq, 0), q_group(nq, -1);
    GroupDesc g;
    std::vector<uint32_t> excl_n(32, 0);
    auto reset = [&]() {
        memset(&g, 0, sizeof(g));
        for (int l = 0; l < 32; ++l) { g.row[l] = -1; g.qidx[l] = -1; }
        std::fill(excl_n.begin(), excl_n.end(), 0u);
        g.adj_begin[0] = (int32_t)plan.excl.size();
    };
    reset();
    const uint32_t budget = (uint32_t)s->max_smem - scan_fixed_smem(Kp) - 256u;
    const uint32_t lbudget = lean ? (uint32_t)s->max_smem - lean_fixed_smem(Ks) - 256u : 0u;
    auto close_group = [&]() {
        if (g.n_lanes == 0) return;
        for (int l = g.n_lanes; l < 32; ++l) g.adj_begin[l + 1] = g.adj_begin[g.n_lanes];
        finish_group(s, plan, g, excl_n.data(), Kp, lean, Ks);
        reset();
    };
    for (int q = 0; q < nq; ++q) {
        const int r = s->row_of(q_uid[q]);
        if (r < 0) continue;
        const int32_t *ab = nullptr, *ae = nullptr;
        uint32_t nex = 0;
        if (exclude_mode >= PK_EXCLUDE_SELF_AND_ADJ && s->adj_of(q_uid[q], &ab, &ae)) nex = (uint32_t)(ae - ab);
        if (g.n_lanes > 0) {   // would adding this query overflow a shared-memory budget?
            GroupDesc t = g;
            t.row[t.n_lanes] = r;
            std::vector<uint32_t> en = excl_n;
            en[t.n_lanes] = nex;
            ++t.n_lanes;
            layout_group(s, t, en.data(), lean);
            if (t.arena_bytes - t.exact_begin > budget || (lean && t.lean_end > lbudget)) close_group();
        }
        const int l = g.n_lanes;
        g.row[l] = r; g.qidx[l] = q; excl_n[l] = nex;
        if (nex) plan.excl.insert(plan.excl.end(), ab, ae);
        g.adj_begin[l + 1] = (int32_t)plan.excl.size();
        q_lane[q] = l;
        q_group[q] = (int32_t)plan.groups.size();
        q_nl[q] = 1;
        ++g.n_lanes;
        if (g.n_lanes == 32) close_group();
    }
    close_group();
    const int ng = (int)plan.groups.size();
    J.n_groups = ng;
    if (ng == 0) { J.valid = true; J.gen = s->gen; return PK_OK; }
    uint64_t off = 0;
    for (auto& gg : plan.groups) {
        gg.arena_off = (uint32_t)off;
        off += gg.arena_bytes;
        if (off > 0xF0000000ull) return fail(PK_ERR_UNSUPPORTED, "query batch too large for one arena (split the batch)");
    }
    J.arena_bytes = off;
    J.max_smem_arena = plan.max_smem_arena;
    J.max_smem_lean = plan.max_smem_lean;
    J.with_excl = exclude_mode >= PK_EXCLUDE_SELF_AND_ADJ;
    TRY(J.arena.reserve(std::max<uint64_t>(off, 16)));
    TRY(upload(s, J.groups, plan.groups));
    if (J.with_excl) TRY(upload(s, J.excl, plan.excl));

    // exact work items: every group scans the whole shard, split into slices so that the grid fills the GPU
    // (the whole scan without the pre-filter; with it, only the CTAs of groups that failed the shortlist check run)
    int sm_count = 148;
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, s->device);
    // one CTA per SM (shared memory holds one group's tables): the grid runs in waves of sm_count CTAs, so pick the
    // split count (>= ~6 waves for balance) whose last wave is fullest -- 128 groups x 7 slices = 896 CTAs would
    // leave the 7th wave 5 % full and cost 13 % of the launch, 128 x 8 = 1024 fills 6.92 waves
    int64_t splits = 1;
    {
        const int64_t lo = std::max<int64_t>(1, ((int64_t)sm_count * 6 + ng - 1) / ng), hi = 2 * lo + 1;
        double best = -1.0;
        for (int64_t sp = lo; sp <= hi; ++sp) {
            const int64_t ctas = sp * ng, waves = (ctas + sm_count - 1) / sm_count;
            const double eff = (double)ctas / (double)(waves * sm_count);
            if (eff > best + 1e-9) { best = eff; splits = sp; }
        }
    }
    splits = std::max<int64_t>(1, std::min<int64_t>(splits, (n_cand + 1023) / 1024));
    if (splits > 128 * 8) splits = 128 * 8;
    std::vector<WorkItem> items;
    std::vector<int32_t> g_slot0(ng);
    for (int gi = 0; gi < ng; ++gi) {
        g_slot0[gi] = (int32_t)items.size();
        for (int64_t sp = 0; sp < splits; ++sp) {
            WorkItem it;
            it.group = gi;
            it.cand_begin = (int32_t)(s->shard_begin + n_cand * sp / splits);
            it.cand_end = (int32_t)(s->shard_begin + n_cand * (sp + 1) / splits);
            it.out_slot = (int32_t)items.size();
            items.push_back(it);
        }
        J.pairs += (int64_t)plan.groups[gi].n_lanes * n_cand;
    }
    for (int q = 0; q < nq; ++q) {
        if (q_nl[q] == 0) continue;
        q_slot0[q] = g_slot0[q_group[q]];
        q_nl[q] = (int32_t)splits;
    }
    TRY(J.keys_a.reserve(items.size() * 32 * (size_t)Kp));
    TRY(upload_items(s, plan, items, J.items, &J.n_sm, &J.n_gl));
    TRY(upload(s, J.q_slot0, q_slot0));
    TRY(upload(s, J.q_nl, q_nl));
    TRY(upload(s, J.q_lane, q_lane));

    if (lean) {
        J.lean = true; J.Ks = Ks; J.n_slices = (int)splits; J.sm_count = sm_count;
        J.cand_begin = s->shard_begin; J.cand_end = s->shard_end;
        // pre-filter items (same slicing), groups whose lean tables do not fit shared memory go straight to the
        // exact full scan (flag preset)
        std::vector<WorkItem> pre;
        std::vector<int32_t> pre_slot0(nq, 0), pre_nl(nq, 0), gflag0(ng, 0), qflag0(nq, 0), gp0(ng, 0);
        std::vector<int32_t> dyn_groups;
        for (int gi = 0; gi < ng; ++gi) {
            gp0[gi] = (int32_t)pre.size();
            if (!plan.groups[gi].lean_in_smem) { gflag0[gi] = 1; continue; }
            dyn_groups.push_back(gi);
            for (int64_t sp = 0; sp < splits; ++sp) {
                WorkItem it = items[(size_t)g_slot0[gi] + sp];
                it.out_slot = (int32_t)pre.size();
                pre.push_back(it);
            }
        }
        J.n_dyn = (int)dyn_groups.size();
        TRY(upload(s, J.dyn_groups, dyn_groups));
        TRY(J.dyn_state.reserve(std::max<size_t>(2 * dyn_groups.size(), 1)));      // cursor, instances per group
        if (dyn_groups.size() * 32 > pk_store::kThetaCap) return fail(PK_ERR_UNSUPPORTED, "more than 8192 query groups in one full-scan batch (split the batch)");
        TRY(s->d_theta_x.reserve(2 * pk_store::kThetaCap));
        // exact re-score items: the 32 x Ks shortlist rows of each group, in list mode
        int rr_splits = 4;      // measured: 8 splits (6.92 waves) cost 4.7 ms against 3.9 ms (table staging per CTA)
        if (const char* e = getenv("POKEC_B200_RR_SPLITS")) rr_splits = std::max(1, atoi(e));
        std::vector<WorkItem> rr;
        std::vector<int32_t> rr_slot0(nq, 0), rr_nl(nq, 0), gr0(ng, 0);
        for (int gi = 0; gi < ng; ++gi) {
            gr0[gi] = (int32_t)rr.size();
            if (gflag0[gi]) continue;
            const int64_t base = (int64_t)gi * 32 * Ks, len = (int64_t)32 * Ks;
            for (int sp = 0; sp < rr_splits; ++sp) {
                WorkItem it;
                it.group = gi;
                it.cand_begin = (int32_t)(base + len * sp / rr_splits);
                it.cand_end = (int32_t)(base + len * (sp + 1) / rr_splits);
                it.out_slot = (int32_t)rr.size();
                rr.push_back(it);
            }
            J.rr_pairs += (int64_t)plan.groups[gi].n_lanes * len;
        }
        for (int q = 0; q < nq; ++q) {
            const int gi = q_group[q];
            if (gi < 0) continue;
            if (gflag0[gi]) { qflag0[q] = 1; continue; }
            pre_slot0[q] = gp0[gi]; pre_nl[q] = (int32_t)splits;
            rr_slot0[q] = gr0[gi]; rr_nl[q] = rr_splits;
        }
        J.n_pre = pre.size();
        TRY(upload(s, J.pre_items, pre));
        TRY(upload_items(s, plan, rr, J.rr_items, &J.n_rr_sm, &J.n_rr_gl));
        TRY(upload(s, J.pre_slot0, pre_slot0)); TRY(upload(s, J.pre_nl, pre_nl));
        TRY(upload(s, J.rr_slot0, rr_slot0)); TRY(upload(s, J.rr_nl, rr_nl));
        TRY(upload(s, J.q_group, q_group));
        TRY(upload(s, J.q_flag0, qflag0)); TRY(upload(s, J.group_flag0, gflag0));
        TRY(J.q_flag.reserve(nq)); TRY(J.group_flag.reserve(ng));
        TRY(J.keys_pre.reserve(std::max<size_t>(pre.size(), 1) * 32 * (size_t)Ks));
        TRY(J.keys_short.reserve((size_t)nq * Ks));
        TRY(J.keys_rr.reserve(std::max<size_t>(rr.size(), 1) * 32 * (size_t)Kp));
        J.n_rr = rr.size();
        TRY(J.short_rows.reserve((size_t)ng * 32 * Ks));
        {
            std::vector<int32_t> seg((size_t)nq + 1);
            for (int q = 0; q <= nq; ++q) seg[(size_t)q] = q * Ks;
            TRY(upload(s, J.rs_seg, seg));
            TRY(J.rs_a.reserve((size_t)nq * Ks)); TRY(J.rs_b.reserve((size_t)nq * Ks)); TRY(J.rs_id.reserve((size_t)nq * Ks));
            TRY(J.rs_f.reserve((size_t)nq * Ks));
            TRY(J.keys_rr.reserve((size_t)nq * Ks));
        }
        {   // Scan order (PreParams): the 1/16 of the shard's rows with the most tokens come first (measured on the
            // bench world, pre-filter ms at 1 GPU / on a 1/8 shard: none 85.0 / 14.4, 1/256 81.5 / 13.6, 1/64 79.7 / 13.3, 1/16 78.4 / 13.0)
            int div = 16;
            if (const char* e = getenv("POKEC_B200_SEED_DIV")) div = atoi(e);
            J.n_seed = 0;
            if (div > 0 && n_cand >= 65536) {
                if (s->seed_b != s->shard_begin || s->seed_e != s->shard_end || s->seed_div != div) {
                    const int64_t want = n_cand / div;
                    std::vector<int32_t> rows((size_t)n_cand);
                    for (int64_t i = 0; i < n_cand; ++i) rows[(size_t)i] = (int32_t)(s->shard_begin + i);
                    std::nth_element(rows.begin(), rows.begin() + want, rows.end(),
                                     [&](int32_t a, int32_t b) { return s->h_ntok[a] != s->h_ntok[b] ? s->h_ntok[a] > s->h_ntok[b] : a < b; });
                    rows.resize((size_t)want);
                    std::sort(rows.begin(), rows.end(),
                              [&](int32_t a, int32_t b) { return s->h_ntok[a] != s->h_ntok[b] ? s->h_ntok[a] > s->h_ntok[b] : a < b; });
                    std::vector<uint8_t> skip((size_t)s->n, 0);
                    for (int32_t r : rows) skip[(size_t)r] = 1;
                    TRY(upload(s, s->d_seed_rows, rows));
                    TRY(upload(s, s->d_seed_skip, skip));
                    CK(cudaStreamSynchronize(s->stream));
                    s->seed_cache.swap(rows);
                    s->seed_b = s->shard_begin; s->seed_e = s->shard_end; s->seed_div = div;
                }
                J.n_seed = (int)s->seed_cache.size();
            }
        }
    }
    CK(cudaStreamSynchronize(s->stream));
    J.valid = true;
    J.gen = s->gen;
    return PK_OK;
}

int launch_prefilter(pk_store* s, const PreParams& P, size_t n_items, uint32_t lean_arena, int Ks, bool store) {
    if (n_items == 0) return PK_OK;
    const uint32_t smem = lean_arena + lean_fixed_smem(store ? 0 : Ks);
    if (store) {
        auto kern = fas_prefilter_kernel<true>;
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)n_items, kPreThreads, smem, s->stream>>>(P);
    } else {
        auto kern = fas_prefilter_kernel<false>;
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)n_items, kPreThreads, smem, s->stream>>>(P);
    }
    CK(cudaGetLastError());
    ++s->stat_launches;
    return PK_OK;
}

PreParams base_pre_params(pk_store* s) {
    PreParams P;
    memset(&P, 0, sizeof(P));
    P.lrec = s->d_lrec.p; P.lrec_off = s->d_lrec_off.p; P.lmask = s->d_lmask.p; P.ln = s->d_lnorm.p;
    P.tab_comp = s->d_tab_comp32.p; P.tab_age = s->d_tab_age32.p;
    return P;
}

PreParams job_pre_params(pk_store* s) {
    FullscanJob& J = s->job;
    PreParams Q = base_pre_params(s);
    Q.groups = J.groups.p; Q.items = J.pre_items.p; Q.arena = J.arena.p;
    Q.Ks = J.Ks; Q.exclude_mode = J.exclude; Q.smem_arena_bytes = J.max_smem_lean; Q.out_keys = J.keys_pre.p;
    Q.dyn_cursor = J.dyn_state.p; Q.dyn_ninst = J.dyn_state.p + J.n_dyn; Q.dyn_groups = J.dyn_groups.p;
    Q.n_dyn = J.n_dyn; Q.max_inst = J.n_slices;
    Q.dyn_theta = s->d_theta_x.p;
    Q.n_peers = s->n_peers;
    for (int r = 0; r < 7; ++r) Q.peer_theta[r] = s->peer_theta[r];
    Q.theta_tag = (J.plan_fp ^ (s->exec_counter * 0x9E3779B1u)) | 1u;
    // the second half of the allocation: the rank-k_frac scores (PreParams::dyn_sfrac)
    Q.dyn_sfrac = s->d_theta_x.p + pk_store::kThetaCap;
    for (int r = 0; r < 7; ++r) Q.peer_sfrac[r] = s->peer_theta[r] ? s->peer_theta[r] + pk_store::kThetaCap : nullptr;
    Q.k_frac = (J.K + s->n_peers) / (s->n_peers + 1);
    Q.K = J.K;
    Q.theta_drop = (getenv("POKEC_B200_NO_PRUNE") && getenv("POKEC_B200_NO_PRUNE")[0] == '1') ? 1e9f : 2.0f * s->margin + 1e-5f;
    Q.cand_begin = (int)J.cand_begin; Q.cand_end = (int)J.cand_end;
    if (J.n_seed > 0) { Q.row_list = s->d_seed_rows.p; Q.skip = s->d_seed_skip.p; Q.n_list = J.n_seed; }
    return Q;
}

// query tables (+ the per-execution state of the two-stage scan)
int fullscan_begin(pk_store* s) {
    FullscanJob& J = s->job;
    TRY(job_ok(s));
    s->stat_scan_ns = 0; s->stat_scan_launches = 0; s->stat_prefilter_ns = 0;
    s->stat_pairs = J.pairs; s->stat_groups = J.n_groups;
    CK(cudaMemsetAsync(J.keys_b.p, 0, (size_t)std::max(J.nq, 1) * J.K * 8, s->stream));
    if (J.n_groups == 0) return PK_OK;
    if (J.lean) CK(cudaEventRecord(s->pev[0], s->stream));
    CK(cudaMemsetAsync(J.arena.p, 0, std::max<uint64_t>(J.arena_bytes, 16), s->stream));
    PrepParams PP;
    PP.rec = s->d_rec.p; PP.rec_off = s->d_rec_off.p; PP.groups = J.groups.p; PP.arena = J.arena.p;
    PP.excl_ids = J.with_excl ? J.excl.p : nullptr;
    PP.lean_fail = nullptr;
    if (J.lean) {
        TRY(J.lean_fail.reserve((size_t)J.n_groups));
        CK(cudaMemsetAsync(J.lean_fail.p, 0, (size_t)J.n_groups * 4, s->stream));
        PP.lean_fail = J.lean_fail.p;
    }
    group_prep_kernel<<<(unsigned)J.n_groups, kPrepThreads, 0, s->stream>>>(PP);
    ++s->stat_launches;
    CK(cudaGetLastError());
    if (!J.lean) return PK_OK;
    const int Ks = J.Ks;
    CK(cudaMemcpyAsync(J.q_flag.p, J.q_flag0.p, (size_t)J.nq * 4, cudaMemcpyDeviceToDevice, s->stream));
    CK(cudaMemcpyAsync(J.group_flag.p, J.group_flag0.p, (size_t)J.n_groups * 4, cudaMemcpyDeviceToDevice, s->stream));
    CK(cudaMemsetAsync(J.short_rows.p, 0xFF, (size_t)J.n_groups * 32 * Ks * 4, s->stream));
    CK(cudaMemsetAsync(J.keys_short.p, 0, (size_t)J.nq * Ks * 8, s->stream));
    // cursors, instance counters and thresholds start at 0; a new execution, a new threshold tag
    CK(cudaMemsetAsync(J.dyn_state.p, 0, std::max<size_t>(2 * (size_t)J.n_dyn, 1) * 4, s->stream));
    CK(cudaMemsetAsync(s->d_theta_x.p, 0, std::max<size_t>((size_t)J.n_dyn * 32, 1) * 8, s->stream));
    CK(cudaMemsetAsync(s->d_theta_x.p + pk_store::kThetaCap, 0, std::max<size_t>((size_t)J.n_dyn * 32, 1) * 8, s->stream));
    ++s->exec_counter;
    CK(cudaEventRecord(s->pev[1], s->stream));
    CK(cudaEventRecord(s->ev0, s->stream));
    return PK_OK;
}

int fullscan_shortlists(pk_store* s) {
    FullscanJob& J = s->job;
    TRY(job_ok(s));
    TRY(fullscan_begin(s));
    if (J.n_groups == 0) return PK_OK;
    MergeParams M;
    memset(&M, 0, sizeof(M));
    M.lane = J.q_lane.p; M.nq = J.nq; M.K = J.K; M.out_keys = J.keys_b.p;
    if (!J.lean) {   // exact arithmetic on every pair
        ScanParams P = base_scan_params(s, J.K);
        P.groups = J.groups.p; P.arena = J.arena.p;
        P.exclude_mode = J.exclude;
        P.out_keys = J.keys_a.p;
        TRY((launch_scan<EPI_TOPK, false>(s, P, J.items.p, J.n_sm, J.n_gl, J.max_smem_arena)));
        M.keys = J.keys_a.p; M.slot0 = J.q_slot0.p; M.n_lists_q = J.q_nl.p; M.stride_k = J.Kp;
        return launch_merge(s, M, J.nq);
    }
    // 1. pre-filter: fp32 scores of the shard (heaviest rows first), Ks best per query and instance
    const int Ks = J.Ks;
    PreParams Q = job_pre_params(s);
    // shortlist instances a CTA never takes stay empty; cursors and instance counters restart at 0
    CK(cudaMemsetAsync(J.keys_pre.p, 0, std::max<size_t>(J.n_pre, 1) * 32 * (size_t)Ks * 8, s->stream));
    // one persistent CTA per SM -- minus the SMs a pipelining caller asked to keep free, so that the short kernels and
    // the NCCL all-gathers of the PREVIOUS batch (other stream) find room while this scan runs
    const size_t pre_grid = std::min<size_t>((size_t)std::max(J.sm_count - s->reserve_sms, 1), J.n_pre);
    TRY(launch_prefilter(s, Q, pre_grid, J.max_smem_lean, Ks, false));
    CK(cudaEventRecord(s->ev1, s->stream));
    if (J.n_pre) { s->scan_time_pending = true; ++s->stat_scan_launches; }
    // 2. merge the instances' shortlists per query: the shard's approximate top-Ks
    MergeParams MS;
    memset(&MS, 0, sizeof(MS));
    MS.keys = J.keys_pre.p; MS.slot0 = J.pre_slot0.p; MS.n_lists_q = J.pre_nl.p; MS.lane = J.q_lane.p;
    MS.nq = J.nq; MS.K = Ks; MS.stride_k = Ks; MS.out_keys = J.keys_short.p;
    MS.poison_group_fail = J.lean_fail.p; MS.q_group = J.q_group.p;
    TRY(launch_merge(s, MS, J.nq));
    CK(cudaEventRecord(s->pev[2], s->stream));
    return PK_OK;
}

// Second half of the two-stage scan: proof, exact re-score, fallback, final merge.  With shortlists gathered from
// n_ranks candidate shards (d_all_short = [n_lists][nq][Ks]) the merged list is the approximate top-Ks over ALL
// candidates; every rank proves it for every query (the flags are the same everywhere), re-scores only the groups
// it owns (any rank can: the packed store holds the whole world) and runs the exact fallback of flagged groups
// over its own shard.  The per-rank [nq][k] keys are then merged by the caller exactly as before.
int fullscan_finish(pk_store* s, const uint64_t* d_all_short, int n_lists, int rank, int n_ranks) {
    FullscanJob& J = s->job;
    TRY(job_ok(s));
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks || n_lists < 1) return fail(PK_ERR_ARG, "bad rank / n_ranks / n_lists");
    if (J.n_groups == 0 || !J.lean) return PK_OK;       // the exact scan of the first half already produced the keys
    const int Ks = J.Ks;
    const uint64_t* short_keys = J.keys_short.p;
    CK(cudaEventRecord(s->pev[3], s->stream));
    if (d_all_short) {
        TRY(J.keys_short_all.reserve((size_t)J.nq * Ks));
        MergeParams MG;
        memset(&MG, 0, sizeof(MG));
        MG.keys = d_all_short; MG.n_lists = n_lists; MG.nq = J.nq; MG.K = Ks; MG.stride_k = Ks; MG.out_keys = J.keys_short_all.p;
        TRY(launch_merge(s, MG, J.nq));
        short_keys = J.keys_short_all.p;
    }
    ScanParams P = base_scan_params(s, J.K);
    P.groups = J.groups.p; P.arena = J.arena.p;
    P.exclude_mode = J.exclude;
    MergeParams M;
    memset(&M, 0, sizeof(M));
    M.lane = J.q_lane.p; M.nq = J.nq; M.K = J.K; M.out_keys = J.keys_b.p;
    CheckParams C;
    C.short_keys = short_keys; C.q_group = J.q_group.p; C.q_lane = J.q_lane.p; C.cand_rows = J.short_rows.p;
    C.q_flag = J.q_flag.p; C.group_flag = J.group_flag.p; C.lean_fail = J.lean_fail.p; C.nq = J.nq; C.K = J.K; C.Ks = Ks; C.margin = s->margin;
    // the candidates that can still reach the top-k become (query, candidate) pairs of the lane-per-pair kernel;
    // with several ranks each rank re-scores the query groups it owns (the others' slots stay empty)
    CK(cudaMemsetAsync(J.rs_b.p, 0xFF, (size_t)J.nq * Ks * 4, s->stream));       // queries without a profile write nothing
    C.pa_row = J.rs_a.p; C.pb_row = J.rs_b.p; C.pkey_id = J.rs_id.p; C.uid = s->d_uid.p; C.groups = J.groups.p;
    C.own_lo = (int)((int64_t)J.n_groups * rank / n_ranks); C.own_hi = (int)((int64_t)J.n_groups * (rank + 1) / n_ranks);
    shortlist_check_kernel<<<J.nq, 64, 0, s->stream>>>(C);
    ++s->stat_launches;
    CK(cudaGetLastError());
    // 3. exact arithmetic on the shortlists; 4. exact full scan for the groups with a flagged query
    CK(cudaEventRecord(s->ev2, s->stream));
    TRY((pairs_launch<false>(s, (int64_t)J.nq * Ks, J.rs_a.p, J.rs_b.p, J.rs_f.p, nullptr, J.rs_id.p, J.keys_rr.p, false)));
    ScanParams F = P;
    F.out_keys = J.keys_a.p; F.group_flag = J.group_flag.p;
    TRY((launch_scan_plain<EPI_TOPK, false>(s, F, J.items.p, J.n_sm, J.n_gl, J.max_smem_arena)));
    CK(cudaEventRecord(s->ev3, s->stream));
    s->rescore_time_pending = true;
    // 5. final merge: the re-scored shortlist, or the full exact scan where the shortlist was not proven
    M.keys = J.keys_rr.p; M.seg_begin = J.rs_seg.p;
    M.q_flag = J.q_flag.p; M.alt_keys = J.keys_a.p; M.alt_slot0 = J.q_slot0.p; M.alt_n_lists_q = J.q_nl.p; M.alt_stride_k = J.Kp;
    TRY(launch_merge(s, M, J.nq));
    CK(cudaEventRecord(s->pev[5], s->stream));
    s->phase_pending = J.n_pre > 0;
    return PK_OK;
}

int fullscan_execute(pk_store* s) {
    TRY(fullscan_shortlists(s));
    return fullscan_finish(s, nullptr, 1, 0, 1);
}

int fullscan_fetch(pk_store* s, int32_t* out_id, float* out_score, int32_t* out_cnt) {
    FullscanJob& J = s->job;
    TRY(job_ok(s));
    if (J.nq == 0) return PK_OK;
    if (J.k == 0) { for (int q = 0; q < J.nq; ++q) out_cnt[q] = 0; return PK_OK; }
    std::vector<uint64_t> keys((size_t)J.nq * J.k);
    CK(cudaMemcpyAsync(keys.data(), J.keys_b.p, keys.size() * 8, cudaMemcpyDeviceToHost, s->stream));
    s->stat_d2h += (int64_t)keys.size() * 8;
    std::vector<int32_t> flags;
    if (J.lean && J.n_groups > 0) {
        flags.resize((size_t)J.nq);
        CK(cudaMemcpyAsync(flags.data(), J.q_flag.p, flags.size() * 4, cudaMemcpyDeviceToHost, s->stream));
    }
    CK(cudaStreamSynchronize(s->stream));
    TRY(collect_scan_time(s));
    s->stat_fallback = 0;
    for (int32_t f : flags) s->stat_fallback += f != 0;
    decode_keys(keys.data(), J.nq, J.k, out_id, out_score, out_cnt);
    return PK_OK;
}

// Host threads for the per-user candidate enumeration of a batch (each user is independent; the store is only
// read).  POKEC_B200_HOST_THREADS overrides the count (1 = the calling thread only).
template <class F>
void parallel_for(int n, F fn) {
    int nt = (int)std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 32u);
    if (const char* e = getenv("POKEC_B200_HOST_THREADS")) nt = std::max(1, atoi(e));
    nt = std::min(nt, (n + 15) / 16);
    if (nt <= 1) { for (int i = 0; i < n; ++i) fn(i); return; }
    std::atomic<int> next{0};
    auto work = [&]() {
        for (;;) {
            const int i0 = next.fetch_add(16);
            if (i0 >= n) break;
            for (int i = i0, e = std::min(n, i0 + 16); i < e; ++i) fn(i);
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back(work);
    work();
    for (auto& t : th) t.join();
}

// ---- candidate enumeration on the host (integer, order-sensitive; mirrors the reference) -----------

// src/recommender_graph.cpp:10-31
void gather_candidates(const pk_store* s, int32_t user, int limit, std::vector<int32_t>& out) {
    out.clear();
    const int32_t *b, *e;
    if (!s->adj_of(user, &b, &e)) return;
    std::unordered_set<int32_t> seen;
    for (const int32_t* pf = b; pf != e; ++pf) {
        const int32_t f = *pf;
        if (f == user) continue;
        if (seen.insert(f).second) out.push_back(f);
        if ((int)out.size() >= limit) return;
        const int32_t *b2, *e2;
        if (!s->adj_of(f, &b2, &e2)) continue;
        for (const int32_t* pff = b2; pff != e2; ++pff) {
            const int32_t ff = *pff;
            if (ff == user) continue;
            if (seen.insert(ff).second) {
                out.push_back(ff);
                if ((int)out.size() >= limit) return;
            }
        }
    }
}

// src/recommender_graph.cpp:114-125 (limit < 0 = unlimited, used by the clubs path)
void collab_candidates(const pk_store* s, int32_t user, const int32_t* fb, const int32_t* fe, int limit, bool limited,
                       std::vector<int32_t>& out) {
    out.clear();
    std::unordered_set<int32_t> seen;
    for (const int32_t* pf = fb; pf != fe; ++pf) {
        const int32_t *b2, *e2;
        if (!s->adj_of(*pf, &b2, &e2)) continue;
        for (const int32_t* p = b2; p != e2; ++p) {
            const int32_t fof = *p;
            if (fof == user) continue;
            if (seen.insert(fof).second) out.push_back(fof);
            if (limited && (int)out.size() >= limit) break;
        }
        if (limited && (int)out.size() >= limit) break;
    }
}

GroupDesc empty_group() {
    GroupDesc g;
    memset(&g, 0, sizeof(g));
    for (int l = 0; l < 32; ++l) { g.row[l] = -1; g.qidx[l] = -1; }
    return g;
}

// ---- recommend: graph / interest ---------------------------------------------------------------------
// recommend_graph_registration == recommend_by_interest (src/recommender_graph.cpp:33-103,224-227): every
// (query, candidate) of the batch is one pair of the lane-per-pair kernel, which also emits the sort keys; the
// sort-based selection then takes each query's top-k from its slice (any k: src/test.cpp:75 passes |friends| / 5).
int recommend_graph(pk_store* s, int nq, const int32_t* q_uid, int k, int limit, uint64_t* h_keys) {
    const int K = std::max(k, 1);
    // phase 1 (host threads): the candidate rows of every query, in the reference's order
    std::vector<std::vector<int32_t>> q_rows((size_t)nq);
    std::vector<int32_t> q_row((size_t)nq, -1);
    parallel_for(nq, [&](int q) {
        const int32_t user = q_uid[q];
        const int r = s->row_of(user);
        q_row[q] = r;
        if (r < 0) return;                                      // src/recommender_graph.cpp:40
        std::vector<int32_t> cands;
        gather_candidates(s, user, limit, cands);
        std::unordered_set<int32_t> existing;                   // :46-49
        const int32_t *b, *e;
        if (s->adj_of(user, &b, &e)) existing.insert(b, e);
        existing.insert(user);
        std::vector<int32_t>& out = q_rows[q];
        for (int32_t c : cands) {
            if (existing.count(c)) continue;
            const int rc = s->row_of(c);
            if (rc < 0) continue;                               // :54
            out.push_back(rc);
        }
    });
    // phase 2: flat pair arrays, one slice per query
    std::vector<int32_t> q_begin((size_t)nq + 1, 0);
    for (int q = 0; q < nq; ++q) q_begin[q + 1] = q_begin[q] + (int32_t)q_rows[q].size();
    const int64_t np = q_begin[nq];
    if (np > 0x7fffff00ll) return fail(PK_ERR_UNSUPPORTED, "graph batch too large (split the batch)");
    std::vector<int32_t> a_row((size_t)np), b_row((size_t)np), b_uid((size_t)np);
    parallel_for(nq, [&](int q) {
        int32_t o = q_begin[q];
        for (int32_t rc : q_rows[q]) { a_row[(size_t)o] = q_row[q]; b_row[(size_t)o] = rc; b_uid[(size_t)o] = s->uid[(size_t)rc]; ++o; }
    });
    s->stat_pairs += np;
    s->stat_groups = 0;
    std::vector<uint64_t> keys((size_t)std::max(nq, 1) * K, 0ull);
    if (np > 0) {
        TRY(upload(s, s->d_i32a, a_row));
        TRY(upload(s, s->d_i32b, b_row));
        TRY(upload(s, s->d_i32c, b_uid));
        TRY(upload(s, s->d_i32d, q_begin));
        TRY(s->d_f32a.reserve((size_t)np));
        TRY(s->d_keys_a.reserve((size_t)np));
        TRY(s->d_keys_b.reserve((size_t)nq * K));
        TRY((pairs_launch<false>(s, np, s->d_i32a.p, s->d_i32b.p, s->d_f32a.p, nullptr, s->d_i32c.p, s->d_keys_a.p)));
        TRY(launch_slice_topk(s, s->d_keys_a.p, s->d_i32d.p, nq, K, s->d_keys_b.p));
        CK(cudaMemcpyAsync(keys.data(), s->d_keys_b.p, (size_t)nq * K * 8, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        TRY(collect_scan_time(s));
    }
    memcpy(h_keys, keys.data(), (size_t)nq * K * 8);
    return PK_OK;
}

// ---- batched list modes with the candidate expansion on the device (pk_expand.cuh) ---------------------
int ensure_graph(pk_store* s) {
    if (!s->graph_dirty) return PK_OK;
    TRY(upload(s, s->g_adj_ptr, s->adj_ptr));
    TRY(upload(s, s->g_adj, s->adj));
    if (!s->row_direct.empty()) {
        TRY(upload(s, s->g_row_direct, s->row_direct));
        TRY(upload(s, s->g_adj_direct, s->adj_direct));
    } else {
        std::vector<int32_t> ku(s->uid_sorted.size()), kr(s->uid_sorted.size());
        for (size_t i = 0; i < ku.size(); ++i) { ku[i] = s->uid_sorted[i].first; kr[i] = s->uid_sorted[i].second; }
        TRY(upload(s, s->g_uid_s, ku)); TRY(upload(s, s->g_uid_srow, kr));
        std::vector<std::pair<int32_t, int32_t>> a(s->adj_uid.size());
        for (size_t r = 0; r < a.size(); ++r) a[r] = {s->adj_uid[r], (int32_t)r};
        std::sort(a.begin(), a.end());
        std::vector<int32_t> au(a.size()), ar(a.size());
        for (size_t i = 0; i < a.size(); ++i) { au[i] = a[i].first; ar[i] = a[i].second; }
        TRY(upload(s, s->g_adj_suid, au)); TRY(upload(s, s->g_adj_srow, ar));
        CK(cudaStreamSynchronize(s->stream));       // locals
    }
    const size_t E = std::max<size_t>(s->adj.size(), 1);
    TRY(s->g_fr_slot.reserve(E)); TRY(s->g_slot_row.reserve(E));
    TRY(upload(s, s->g_ncl, s->h_ncl));
    {
        std::vector<uint32_t> cost((size_t)s->n);
        for (int r = 0; r < s->n; ++r) cost[(size_t)r] = s->h_ntok[r] + s->h_nfr[r] + s->h_ncl[r];
        TRY(upload(s, s->g_cost, cost));
        TRY(upload(s, s->g_nfr, s->h_nfr));
        TRY(upload(s, s->g_ntok, s->h_ntok));
        CK(cudaStreamSynchronize(s->stream));
    }
    CK(cudaStreamSynchronize(s->stream));
    s->graph_dirty = false;
    return PK_OK;
}

DevGraph dev_graph(pk_store* s) {
    DevGraph G;
    memset(&G, 0, sizeof(G));
    G.adj_ptr = s->g_adj_ptr.p; G.adj = s->g_adj.p; G.uid = s->d_uid.p;
    if (!s->row_direct.empty()) {
        G.adj_direct = s->g_adj_direct.p; G.row_direct = s->g_row_direct.p;
        G.uid_lo = s->uid_lo; G.range = (int64_t)s->row_direct.size();
    } else {
        G.adj_sorted_uid = s->g_adj_suid.p; G.adj_sorted_row = s->g_adj_srow.p;
        G.uid_sorted = s->g_uid_s.p; G.uid_sorted_row = s->g_uid_srow.p;
    }
    G.m = (int)s->adj_uid.size(); G.n = s->n;
    return G;
}

// the device pipeline takes a batch when nothing it mirrors was overridden behind the store and the per-warp scratch
// set stays small; single-user drop-in calls keep the host walk (they must read the caller's LIVE adjacency)
bool device_lists_usable(pk_store* s, int mode, int nq, int limit) {
    if (const char* e = getenv("POKEC_B200_HOST_LISTS")) if (e[0] == '1') return false;
    if (!s->adj_override.empty() || nq < 16) return false;
    if (limit < 1 || limit > 16384 || s->max_deg > 32768) return false;
    if (s->adj.size() > 0x7fffff00ull) return false;
    return true;
}

// expansion of a batch: fills g_qrow / g_cand / g_cnt (+ friend slots for the collaborative mode)
int expand_batch(pk_store* s, int mode, int nq, const int32_t* q_uid, int limit, bool raw, bool slots_only = false) {
    TRY(ensure_graph(s));
    uint32_t cap = 1024;
    while (cap < 2u * (uint32_t)(limit + 32) || cap < 2u * (uint32_t)(s->max_deg + 32)) cap <<= 1;
    int sm_count = 148;
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, s->device);
    const int ctas = std::max(1, std::min(sm_count * 2, (nq + kExpWarps - 1) / kExpWarps));
    const int warps = sm_count * 2 * kExpWarps;
    if (cap != s->g_tab_cap || warps != s->g_warps) {     // new table geometry: every tag is meaningless
        TRY(s->g_scratch.reserve((size_t)warps * cap));
        TRY(s->g_epoch.reserve((size_t)warps));
        CK(cudaMemsetAsync(s->g_scratch.p, 0, (size_t)warps * cap * 8, s->stream));
        CK(cudaMemsetAsync(s->g_epoch.p, 0, (size_t)warps * 4, s->stream));
        s->g_tab_cap = cap; s->g_warps = warps;
    }
    TRY(upload(s, s->d_i32e, q_uid, (size_t)nq));
    TRY(s->g_qrow.reserve(nq)); TRY(s->g_cnt.reserve(nq)); TRY(s->g_nslots.reserve(nq)); TRY(s->g_qoff.reserve(nq));
    TRY(s->g_qdeg.reserve(nq)); TRY(s->g_next.reserve(4));
    TRY(s->g_cand.reserve((size_t)nq * limit));
    CK(cudaMemsetAsync(s->g_next.p, 0, 16, s->stream));
    ExpandParams X;
    memset(&X, 0, sizeof(X));
    X.G = dev_graph(s);
    X.q_uid = s->d_i32e.p; X.nq = nq; X.mode = mode; X.limit = limit; X.raw = raw ? 1 : 0; X.slots_only = slots_only ? 1 : 0;
    X.scratch = s->g_scratch.p; X.epoch = s->g_epoch.p; X.tab_cap = cap; X.next_query = s->g_next.p; X.cand_cap = limit;
    X.q_row = s->g_qrow.p; X.cand_out = s->g_cand.p; X.cand_cnt = s->g_cnt.p;
    if (mode != PK_MODE_GRAPH) {
        X.fr_slot = s->g_fr_slot.p; X.slot_row = s->g_slot_row.p; X.n_slots = s->g_nslots.p;
        X.q_adj_off = s->g_qoff.p; X.q_deg = s->g_qdeg.p;
    }
    expand_kernel<<<ctas, kExpWarps * 32, 0, s->stream>>>(X);
    ++s->stat_launches;
    CK(cudaGetLastError());
    return PK_OK;
}

int recommend_lists_device(pk_store* s, int mode, int nq, const int32_t* q_uid, int k, int limit, uint64_t* h_keys) {
    const int K = std::max(k, 1);
    std::vector<uint64_t> keys((size_t)nq * K, 0ull);
    // batches bounded so that the [nq][limit] candidate block and the pair arrays stay moderate
    const int kQ = 8192;
    for (int q0 = 0; q0 < nq; q0 += kQ) {
        const int m = std::min(kQ, nq - q0);
        TRY(expand_batch(s, mode, m, q_uid + q0, limit, false));
        const bool collab = mode != PK_MODE_GRAPH;
        // the FAS(f, c) matrix: groups of 32 friends through fas_scan_kernel (planned on the device), or -- A/B switch --
        // one lane per (f, c) pair
        const char* pm = getenv("POKEC_B200_COLLAB_PAIRS");
        const bool pair_matrix = collab && pm && pm[0] == '1';
        TRY(s->g_cand_off.reserve((size_t)m + 1)); TRY(s->g_pair_off.reserve((size_t)m + 1));
        offsets_kernel<<<1, 1024, 0, s->stream>>>(m, s->g_cnt.p, collab ? s->g_nslots.p : nullptr, s->g_cand_off.p, s->g_pair_off.p, pair_matrix ? 1 : 0);
        ++s->stat_launches;
        int32_t n_cands = 0;
        int64_t n_pairs = 0;
        CK(cudaMemcpyAsync(&n_cands, s->g_cand_off.p + m, 4, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaMemcpyAsync(&n_pairs, s->g_pair_off.p + m, 8, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        if (n_pairs > 0x7fffff00ll) return fail(PK_ERR_UNSUPPORTED, "list-mode batch too large (%lld pairs; split the batch)", (long long)n_pairs);
        s->stat_pairs += n_pairs;
        if (n_cands == 0) continue;
        TRY(s->d_i32a.reserve((size_t)std::max<int64_t>(n_pairs, 1))); TRY(s->d_i32b.reserve((size_t)std::max<int64_t>(n_pairs, 1)));
        TRY(s->d_i32c.reserve((size_t)std::max<int64_t>(collab ? n_cands : n_pairs, 1)));
        TRY(s->d_i32d.reserve((size_t)n_cands));
        TRY(s->d_f32a.reserve((size_t)std::max<int64_t>(n_pairs, 1)));
        TRY(s->d_keys_a.reserve((size_t)n_cands));
        TRY(s->d_keys_b.reserve((size_t)m * K));
        EmitParams E;
        memset(&E, 0, sizeof(E));
        E.nq = m; E.mode = mode; E.cand_cap = limit; E.q_row = s->g_qrow.p; E.cand_out = s->g_cand.p; E.cand_cnt = s->g_cnt.p;
        E.cand_off = s->g_cand_off.p; E.pair_off = s->g_pair_off.p; E.n_slots = s->g_nslots.p; E.q_adj_off = s->g_qoff.p;
        E.slot_row = s->g_slot_row.p; E.uid = s->d_uid.p; E.a_row = s->d_i32a.p; E.b_row = s->d_i32b.p; E.key_id = s->d_i32c.p;
        E.cand_row_flat = s->d_i32d.p; E.cost = s->g_cost.p;
        const char* lay = getenv("POKEC_B200_COLLAB_LAYOUT");
        E.matrix_pairs = pair_matrix ? 1 : 0;
        E.friend_major = (lay && lay[0] == 'f') ? 1 : 0;      // measured: candidate-major 28.9 ms, friend-major 32.0 ms (72.9 M pairs)
        emit_pairs_kernel<<<m, 256, 0, s->stream>>>(E);
        ++s->stat_launches;
        CK(cudaGetLastError());
        if (!collab) {
            TRY((pairs_launch<false>(s, n_pairs, s->d_i32a.p, s->d_i32b.p, s->d_f32a.p, nullptr, s->d_i32c.p, s->d_keys_a.p)));
        } else {
            TRY((pairs_launch<false>(s, n_pairs, s->d_i32a.p, s->d_i32b.p, s->d_f32a.p, nullptr, nullptr, nullptr)));
            if (pair_matrix) {
                CollabReduce2Params R;
                R.nq = m; R.fas = s->d_f32a.p; R.pair_off = s->g_pair_off.p; R.cand_off = s->g_cand_off.p; R.n_slots = s->g_nslots.p;
                R.q_adj_off = s->g_qoff.p; R.q_deg = s->g_qdeg.p; R.fr_slot = s->g_fr_slot.p; R.key_id = s->d_i32c.p; R.out_keys = s->d_keys_a.p;
                R.friend_major = E.friend_major;
                collab_reduce2_kernel<<<dim3(8, m), 128, 0, s->stream>>>(R);
                ++s->stat_launches;
                CK(cudaGetLastError());
            } else {
                // group planning on the device, then the exact scan kernel in store mode over the candidates of each query
                TRY(s->g_group_off.reserve((size_t)m + 1)); TRY(s->g_gtotals.reserve(8));
                CollabGroupsParams G;
                memset(&G, 0, sizeof(G));
                G.nq = m; G.T = s->T; G.smem_budget = (uint32_t)s->max_smem - scan_fixed_smem(0) - 256u;
                G.cand_cnt = s->g_cnt.p; G.cand_off = s->g_cand_off.p; G.n_slots = s->g_nslots.p; G.q_adj_off = s->g_qoff.p;
                G.slot_row = s->g_slot_row.p; G.ncl = s->g_ncl.p; G.nfr = s->g_nfr.p; G.ntok = s->g_ntok.p;
                G.group_off = s->g_group_off.p; G.totals = s->g_gtotals.p;
                collab_groups_count_kernel<<<1, 1024, 0, s->stream>>>(G);
                unsigned long long tot[5] = {0, 0, 0, 0, 0};
                CK(cudaMemcpyAsync(tot, s->g_gtotals.p, 8, cudaMemcpyDeviceToHost, s->stream));
                CK(cudaStreamSynchronize(s->stream));
                const size_t ng = (size_t)tot[0];
                TRY(s->d_groups.reserve(std::max<size_t>(ng, 1))); TRY(s->d_items.reserve(std::max<size_t>(ng, 1)));
                G.groups = s->d_groups.p; G.items = s->d_items.p;
                collab_groups_fill_kernel<<<1, 1024, 0, s->stream>>>(G);
                s->stat_launches += 2;
                CK(cudaMemcpyAsync(tot, s->g_gtotals.p, 40, cudaMemcpyDeviceToHost, s->stream));
                CK(cudaStreamSynchronize(s->stream));
                CK(cudaGetLastError());
                const unsigned long long arena_bytes = tot[1], out_rows = tot[2];
                if (arena_bytes > 0xF0000000ull || out_rows > 0x7fffff00ull / 32)
                    return fail(PK_ERR_UNSUPPORTED, "collaborative batch too large (split the batch)");
                s->stat_groups += (int64_t)ng;
                if (ng > 0) {
                    TRY(s->d_arena.reserve(std::max<uint64_t>(arena_bytes, 16)));
                    CK(cudaMemsetAsync(s->d_arena.p, 0, std::max<uint64_t>(arena_bytes, 16), s->stream));
                    TRY(s->d_f32b.reserve((size_t)std::max<unsigned long long>(out_rows, 1) * 32));
                    PrepParams PP;
                    PP.rec = s->d_rec.p; PP.rec_off = s->d_rec_off.p; PP.groups = s->d_groups.p; PP.arena = s->d_arena.p;
                    PP.excl_ids = nullptr; PP.lean_fail = nullptr;
                    group_prep_kernel<<<(unsigned)ng, kPrepThreads, 0, s->stream>>>(PP);
                    ++s->stat_launches;
                    CK(cudaGetLastError());
                    ScanParams SP = base_scan_params(s, 1);
                    SP.cand_rows = s->d_i32d.p;                        // the flat candidate rows of the batch
                    SP.out_f = s->d_f32b.p;
                    const uint32_t fixed = scan_fixed_smem(0);
                    CK(cudaEventRecord(s->ev2, s->stream));
                    if (ng > (size_t)tot[4]) {
                        SP.want_smem = 1; SP.smem_arena_bytes = (uint32_t)tot[3];
                        TRY((launch_scan_t<EPI_STORE, false>(s, SP, s->d_items.p, ng, true, (uint32_t)tot[3] + fixed)));
                    }
                    if (tot[4] > 0) {
                        SP.want_smem = 0; SP.smem_arena_bytes = 0;
                        TRY((launch_scan_t<EPI_STORE, false>(s, SP, s->d_items.p, ng, false, fixed)));
                    }
                    CK(cudaEventRecord(s->ev3, s->stream));
                    s->rescore_time_pending = false;
                }
                CollabReduce3Params R;
                R.nq = m; R.w = s->d_f32a.p; R.mat = s->d_f32b.p; R.pair_off = s->g_pair_off.p; R.cand_off = s->g_cand_off.p;
                R.n_slots = s->g_nslots.p; R.q_adj_off = s->g_qoff.p; R.q_deg = s->g_qdeg.p; R.fr_slot = s->g_fr_slot.p;
                R.group_off = s->g_group_off.p; R.key_id = s->d_i32c.p; R.items = s->d_items.p; R.out_keys = s->d_keys_a.p;
                collab_reduce3_kernel<<<dim3(8, m), 128, 0, s->stream>>>(R);
                ++s->stat_launches;
                CK(cudaGetLastError());
                // the matrix kernels' time joins the scan-kernel statistic
                CK(cudaStreamSynchronize(s->stream));
                if (ng > 0) {
                    float ms = 0.f;
                    CK(cudaEventElapsedTime(&ms, s->ev2, s->ev3));
                    s->stat_scan_ns += (int64_t)((double)ms * 1e6);
                    s->stat_pairs += (int64_t)out_rows * 32;
                }
            }
        }
        TRY(launch_slice_topk(s, s->d_keys_a.p, s->g_cand_off.p, m, K, s->d_keys_b.p));
        CK(cudaMemcpyAsync(keys.data() + (size_t)q0 * K, s->d_keys_b.p, (size_t)m * K * 8, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        TRY(collect_scan_time(s));
    }
    memcpy(h_keys, keys.data(), (size_t)nq * K * 8);
    return PK_OK;
}

// recommend_clubs_collab for a batch, the edge stream built on the device (clubs_plan_kernel); the FAS values come
// from the per-edge cache when it is valid, else from the lane-per-pair kernel
int recommend_clubs_device(pk_store* s, int nq, const int32_t* q_uid, int k, uint64_t* h_keys) {
    const int K = std::max(k, 1);
    std::vector<uint64_t> keys((size_t)nq * K, 0ull);
    const bool cached = s->edge_cache_valid;
    int sm_count = 148;
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, s->device);
    const int kQ = 8192;
    for (int q0 = 0; q0 < nq; q0 += kQ) {
        const int m = std::min(kQ, nq - q0);
        TRY(expand_batch(s, PK_MODE_CLUBS, m, q_uid + q0, 1, false, true));
        TRY(s->g_nedges.reserve(m)); TRY(s->g_ntwo.reserve(m)); TRY(s->g_ncon.reserve(m));
        TRY(s->g_edge_off.reserve((size_t)m + 1)); TRY(s->g_pair_off.reserve((size_t)m + 1)); TRY(s->g_slot_begin.reserve((size_t)m + 1));
        TRY(s->g_totals.reserve(4));
        ClubsPlanParams C;
        memset(&C, 0, sizeof(C));
        C.G = dev_graph(s); C.q_uid = s->d_i32e.p; C.nq = m;
        C.q_row = s->g_qrow.p; C.n_slots = s->g_nslots.p; C.q_adj_off = s->g_qoff.p; C.q_deg = s->g_qdeg.p;
        C.fr_slot = s->g_fr_slot.p; C.slot_row = s->g_slot_row.p; C.ncl = s->g_ncl.p; C.cached = cached ? 1 : 0;
        C.n_edges = s->g_nedges.p; C.n_two = s->g_ntwo.p; C.n_con = s->g_ncon.p;
        const int ctas = std::max(1, std::min(sm_count * 4, (m + kExpWarps - 1) / kExpWarps));
        clubs_plan_kernel<false><<<ctas, kExpWarps * 32, 0, s->stream>>>(C);
        clubs_offsets_kernel<<<1, 1024, 0, s->stream>>>(m, cached ? 1 : 0, s->g_nedges.p, s->g_ntwo.p, s->g_nslots.p, s->g_ncon.p,
                                                      s->g_edge_off.p, s->g_pair_off.p, s->g_slot_begin.p, s->g_totals.p);
        s->stat_launches += 2;
        int64_t tot[3] = {0, 0, 0};
        CK(cudaMemcpyAsync(tot, s->g_totals.p, 24, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        CK(cudaGetLastError());
        const int64_t n_edges = tot[0], n_pairs = tot[1], n_slots = tot[2];
        if (n_edges > 0x7fffff00ll || n_pairs > 0x7fffff00ll || n_slots > 0x7fffff00ll)
            return fail(PK_ERR_UNSUPPORTED, "club batch too large (split the batch)");
        s->stat_pairs += n_pairs;
        if (n_edges == 0) continue;
        TRY(s->d_edges.reserve((size_t)n_edges));
        TRY(s->d_u32a.reserve((size_t)n_edges));
        TRY(s->d_i32a.reserve((size_t)std::max<int64_t>(n_pairs, 1))); TRY(s->d_i32b.reserve((size_t)std::max<int64_t>(n_pairs, 1)));
        TRY(s->d_f32a.reserve((size_t)std::max<int64_t>(n_pairs, 1)));
        TRY(s->d_i32d.reserve((size_t)n_slots)); TRY(s->d_keys_a.reserve((size_t)n_slots)); TRY(s->d_f64a.reserve((size_t)n_slots));
        TRY(s->d_keys_b.reserve((size_t)m * K));
        C.edge_off = s->g_edge_off.p; C.pair_off = s->g_pair_off.p; C.edges = s->d_edges.p; C.con_off = s->d_u32a.p;
        C.a_row = s->d_i32a.p; C.b_row = s->d_i32b.p;
        clubs_plan_kernel<true><<<ctas, kExpWarps * 32, 0, s->stream>>>(C);
        ++s->stat_launches;
        CK(cudaGetLastError());
        if (!cached) TRY((pairs_launch<false>(s, n_pairs, s->d_i32a.p, s->d_i32b.p, s->d_f32a.p, nullptr, nullptr, nullptr)));
        CK(cudaMemsetAsync(s->d_keys_a.p, 0, (size_t)n_slots * 8, s->stream));
        CK(cudaMemsetAsync(s->d_f64a.p, 0, (size_t)n_slots * 8, s->stream));
        ClubParams P;
        P.rec = s->d_rec.p; P.rec_off = s->d_rec_off.p; P.fas = cached ? s->d_edge_fas.p : s->d_f32a.p; P.edges = s->d_edges.p;
        P.con_off = s->d_u32a.p; P.user_edge_begin = s->g_edge_off.p; P.user_slot_begin = s->g_slot_begin.p;
        P.user_row = s->g_qrow.p; P.con_slot = s->d_i32d.p; P.slot_key = s->d_keys_a.p; P.slot_acc = s->d_f64a.p;
        TRY(s->d_keys_c.reserve((size_t)n_slots)); TRY(s->g_ntwo.reserve((size_t)m));
        P.out_keys = s->d_keys_c.p; P.out_cnt = s->g_ntwo.p;        // (n_two is not read after the FILL pass)
        clubs_scatter_kernel<<<m, kClubThreads, 0, s->stream>>>(P);
        ++s->stat_launches;
        CK(cudaGetLastError());
        TRY(launch_slice_topk(s, s->d_keys_c.p, s->g_slot_begin.p, m, K, s->d_keys_b.p, s->g_ntwo.p));
        CK(cudaMemcpyAsync(keys.data() + (size_t)q0 * K, s->d_keys_b.p, (size_t)m * K * 8, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        TRY(collect_scan_time(s));
    }
    memcpy(h_keys, keys.data(), (size_t)nq * K * 8);
    return PK_OK;
}

// ---- recommend: collaborative friends / clubs --------------------------------------------------------
//
// Both need w_f = FAS(user, f) for the user's friends and the matrix FAS(f, c) of friends x
// friend-of-friend candidates (src/recommender_graph.cpp:127-136,167-180; src/recommender_clubs.cpp:23-64).
// One store-mode scan computes both: per user one 1-lane group {user} x friends, and per 32 distinct
// friends one group x candidates.
struct CollabUser {
    int q = -1;                       // caller-side query index
    int32_t uid = 0;
    int row = -1;
    std::vector<int32_t> friends;     // adjacency row, in order (duplicates kept)
    std::vector<int32_t> fr_slot;     // per friend entry: index among the distinct friends with a profile, -1 = none
    std::vector<int32_t> slot_row;    // row of each distinct friend
    std::vector<int32_t> slot_uid;
    std::vector<int32_t> cands;       // candidate uids with a profile, in reference order
    std::vector<int32_t> cand_row;
    int32_t w_out0 = 0;               // first output row of the {user} x distinct-friends scan
    int32_t mat_out0 = 0;             // first output row of the friends x cands scan ([group][cand])
};

int collab_prepare(pk_store* s, int nq, const int32_t* q_uid, int limit, bool limited, std::vector<CollabUser>& users,
                   Plan& plan, std::vector<WorkItem>& items, std::vector<int32_t>& cand_rows, int64_t* out_rows) {
    int64_t rows = 0;
    // phase 1 (host threads): friends, distinct-friend slots and the ordered candidate list of every user
    std::vector<CollabUser> all((size_t)nq);
    parallel_for(nq, [&](int q) {
        CollabUser& u = all[q];
        u.q = q; u.uid = q_uid[q]; u.row = s->row_of(u.uid);
        if (u.row < 0) return;                                           // src/recommender_graph.cpp:130
        const int32_t *b, *e;
        if (s->adj_of(u.uid, &b, &e)) u.friends.assign(b, e);
        std::unordered_map<int32_t, int32_t> slot_of;
        for (int32_t f : u.friends) {
            const int rf = s->row_of(f);
            if (rf < 0) { u.fr_slot.push_back(-1); continue; }           // :134
            auto ins = slot_of.emplace(f, (int32_t)u.slot_row.size());
            if (ins.second) { u.slot_row.push_back(rf); u.slot_uid.push_back(f); }
            u.fr_slot.push_back(ins.first->second);
        }
        std::vector<int32_t> tmp;
        collab_candidates(s, u.uid, u.friends.data(), u.friends.data() + u.friends.size(), limit, limited, tmp);
        for (int32_t c : tmp) {
            const int rc = s->row_of(c);
            if (rc < 0) continue;                                        // :172
            u.cands.push_back(c);
            u.cand_row.push_back(rc);
        }
    });
    // phase 2: groups and work items, in caller order
    for (int q = 0; q < nq; ++q) {
        CollabUser& u = all[q];
        if (u.row < 0) continue;
        const int nslots = (int)u.slot_row.size();
        if (nslots > 0) {
            // {user} x friends
            GroupDesc g = empty_group();
            g.n_lanes = 1; g.row[0] = u.row; g.qidx[0] = q;
            WorkItem it;
            it.group = (int32_t)plan.groups.size();
            it.cand_begin = (int32_t)cand_rows.size();
            cand_rows.insert(cand_rows.end(), u.slot_row.begin(), u.slot_row.end());
            it.cand_end = (int32_t)cand_rows.size();
            it.out_slot = (int32_t)rows;
            u.w_out0 = (int32_t)rows;
            rows += nslots;
            finish_group(s, plan, g, nullptr, 0);
            items.push_back(it);
            s->stat_pairs += nslots;
            // friends x candidates
            if (!u.cands.empty()) {
                const int32_t cb = (int32_t)cand_rows.size();
                cand_rows.insert(cand_rows.end(), u.cand_row.begin(), u.cand_row.end());
                const int32_t ce = (int32_t)cand_rows.size();
                u.mat_out0 = (int32_t)rows;
                for (int g0 = 0; g0 < nslots; g0 += 32) {
                    GroupDesc gg = empty_group();
                    gg.n_lanes = std::min(32, nslots - g0);
                    for (int l = 0; l < gg.n_lanes; ++l) { gg.row[l] = u.slot_row[g0 + l]; gg.qidx[l] = q; }
                    WorkItem it2;
                    it2.group = (int32_t)plan.groups.size();
                    it2.cand_begin = cb; it2.cand_end = ce; it2.out_slot = (int32_t)rows;
                    rows += ce - cb;
                    finish_group(s, plan, gg, nullptr, 0);
                    items.push_back(it2);
                    s->stat_pairs += (int64_t)gg.n_lanes * (ce - cb);
                }
            }
        }
        if (rows > 0x7fffff00ll / 32) return fail(PK_ERR_UNSUPPORTED, "collaborative batch too large (split the batch)");
        users.push_back(std::move(u));
    }
    *out_rows = rows;
    return PK_OK;
}

int recommend_collab(pk_store* s, int nq, const int32_t* q_uid, int k, int limit, uint64_t* h_keys) {
    const int K = std::max(k, 1);
    std::vector<CollabUser> users;
    Plan plan;
    std::vector<WorkItem> items;
    std::vector<int32_t> cand_rows;
    int64_t rows = 0;
    TRY(collab_prepare(s, nq, q_uid, limit, true, users, plan, items, cand_rows, &rows));
    s->stat_groups = (int64_t)plan.groups.size();
    std::vector<uint64_t> keys((size_t)std::max(nq, 1) * K, 0ull);
    const int nu = (int)users.size();
    if (nu > 0) {
        // a user none of whose friends has a profile still gets its candidates, each with score 0
        // (src/recommender_graph.cpp:167-213: the sum over friends is empty, the candidate is emitted anyway),
        // so the reduce and the top-k run whenever there are candidates; only the scan needs work items
        TRY(s->d_f32a.reserve((size_t)std::max<int64_t>(rows, 1) * 32));
        if (!items.empty()) {
            TRY(upload_plan(s, plan, false));
            TRY(upload(s, s->d_cand_rows, cand_rows));
            ScanParams P = base_scan_params(s, K);
            P.cand_rows = s->d_cand_rows.p;
            P.out_f = s->d_f32a.p;
            TRY((run_scan<EPI_STORE, false>(s, plan, P, items)));
        }

        // reduce in friend order, then top-k per user
        std::vector<int32_t> ucb(nu + 1, 0), ufb(nu + 1, 0), fr_slot, cand_uid, u_ncand(nu);
        std::vector<int64_t> u_w0(nu), u_mat0(nu);
        for (int i = 0; i < nu; ++i) {
            const CollabUser& u = users[i];
            ucb[i + 1] = ucb[i] + (int32_t)u.cands.size();
            ufb[i + 1] = ufb[i] + (int32_t)u.friends.size();
            fr_slot.insert(fr_slot.end(), u.fr_slot.begin(), u.fr_slot.end());
            cand_uid.insert(cand_uid.end(), u.cands.begin(), u.cands.end());
            u_w0[i] = (int64_t)u.w_out0 * 32;
            u_mat0[i] = (int64_t)u.mat_out0 * 32;
            u_ncand[i] = (int32_t)u.cands.size();
        }
        const int total_c = ucb[nu];
        if (total_c > 0) {
            TRY(upload(s, s->d_i32a, ucb));
            TRY(upload(s, s->d_i32b, ufb));
            TRY(upload(s, s->d_i32c, fr_slot));
            TRY(upload(s, s->d_i32d, cand_uid));
            std::vector<int64_t> offs(u_w0);
            offs.insert(offs.end(), u_mat0.begin(), u_mat0.end());
            TRY(upload(s, s->d_i64a, offs));
            TRY(s->d_keys_a.reserve(total_c));
            TRY(s->d_keys_b.reserve((size_t)nu * K));
            CollabReduceParams R;
            R.fas = s->d_f32a.p; R.user_cand_begin = s->d_i32a.p; R.user_fr_begin = s->d_i32b.p; R.fr_slot = s->d_i32c.p;
            R.user_w_off = s->d_i64a.p; R.user_mat_off = s->d_i64a.p + nu; R.cand_uid = s->d_i32d.p;
            R.out_keys = s->d_keys_a.p; R.nu = nu;
            collab_reduce_kernel<<<dim3(8, nu), 128, 0, s->stream>>>(R);
            ++s->stat_launches;
            CK(cudaGetLastError());
            TRY(launch_slice_topk(s, s->d_keys_a.p, s->d_i32a.p, nu, K, s->d_keys_b.p));
            std::vector<uint64_t> uk((size_t)nu * K);
            CK(cudaMemcpyAsync(uk.data(), s->d_keys_b.p, uk.size() * 8, cudaMemcpyDeviceToHost, s->stream));
            CK(cudaStreamSynchronize(s->stream));
            for (int i = 0; i < nu; ++i) memcpy(&keys[(size_t)users[i].q * K], &uk[(size_t)i * K], (size_t)K * 8);
        }
    }
    memcpy(h_keys, keys.data(), (size_t)nq * K * 8);
    return PK_OK;
}

// ---- per-edge FAS cache ----------------------------------------------------------------------------------
// Both weights of recommend_clubs_collab are FAS values on adjacency edges -- w_f = FAS(u, f) for f in adj[u]
// (src/recommender_clubs.cpp:23-28) and s = FAS(f, fof) for fof in adj[f] (:56) -- and do not depend on the query.
// Scored once (the exact kernel, pairs mode), they turn the clubs call into edge enumeration + weighted scatter.
int edge_cache_build(pk_store* s) {
    TRY(ensure_ready(s));
    const int64_t E = (int64_t)s->adj.size();
    // per adjacency entry: (row of the source, row of the neighbour), -1 = no profile -> the kernel stores -1
    std::vector<int32_t> ra((size_t)E), rb((size_t)E);
    const int m = (int)s->adj_uid.size();
    std::atomic<int64_t> valid{0};
    parallel_for((m + 255) / 256, [&](int blk) {
        int64_t v = 0;
        for (int r = blk * 256, e = std::min(m, (blk + 1) * 256); r < e; ++r) {
            const int32_t src = s->row_of(s->adj_uid[(size_t)r]);
            for (int64_t j = s->adj_ptr[(size_t)r]; j < s->adj_ptr[(size_t)r + 1]; ++j) {
                ra[(size_t)j] = src;
                rb[(size_t)j] = s->row_of(s->adj[(size_t)j]);
                v += (src >= 0 && rb[(size_t)j] >= 0) ? 1 : 0;
            }
        }
        valid += v;
    });
    TRY(s->d_edge_fas.reserve((size_t)std::max<int64_t>(E, 1)));
    s->stat_scan_ns = 0; s->stat_scan_launches = 0;
    const int64_t kBatch = 1 << 23;
    for (int64_t p0 = 0; p0 < E; p0 += kBatch) {
        const int64_t n = std::min(kBatch, E - p0);
        TRY(upload(s, s->d_i32a, ra.data() + p0, (size_t)n));
        TRY(upload(s, s->d_i32b, rb.data() + p0, (size_t)n));
        TRY((pairs_launch<false>(s, n, s->d_i32a.p, s->d_i32b.p, s->d_edge_fas.p + p0, nullptr, nullptr, nullptr)));
        CK(cudaStreamSynchronize(s->stream));
        TRY(collect_scan_time(s));
    }
    s->stat_pairs = valid.load();
    s->h_edge_fas.resize((size_t)E);
    if (E > 0) CK(cudaMemcpyAsync(s->h_edge_fas.data(), s->d_edge_fas.p, (size_t)E * 4, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    s->edge_cache_valid = true;
    return PK_OK;
}

int recommend_clubs_cached(pk_store* s, int nq, const int32_t* q_uid, int k, uint64_t* h_keys) {
    const int K = std::max(k, 1);
    std::vector<uint64_t> keys((size_t)std::max(nq, 1) * K, 0ull);
    // users with a profile, in caller order
    std::vector<int> uq; std::vector<int32_t> urow;
    for (int q = 0; q < nq; ++q) {
        const int r = s->row_of(q_uid[q]);
        if (r < 0) continue;                                             // src/recommender_clubs.cpp:16
        uq.push_back(q); urow.push_back(r);
    }
    const int nu = (int)uq.size();
    s->stat_pairs = 0; s->stat_groups = 0; s->stat_scan_ns = 0; s->stat_scan_launches = 0;
    if (nu > 0) {
        // phase 1 (host threads): every user's edge stream in the reference's visiting order -- 1-hop entries, then
        // 2-hop entries; the FAS values are indices into the edge cache (position of the adjacency entry)
        std::vector<std::vector<ClubEdge>> u_edges((size_t)nu);
        std::vector<std::vector<uint32_t>> u_con((size_t)nu);
        std::vector<uint64_t> u_total((size_t)nu, 0);
        parallel_for(nu, [&](int i) {
            const int32_t uid = q_uid[uq[i]];
            std::vector<ClubEdge>& ed = u_edges[i];
            std::vector<uint32_t>& co = u_con[i];
            uint64_t con = 0;
            const int32_t *b, *e;
            if (!s->adj_of(uid, &b, &e)) return;
            const int64_t e0 = b - s->adj.data();
            for (const int32_t* pf = b; pf != e; ++pf) {                 // :34-43
                const int rf = s->row_of(*pf);
                if (rf < 0) continue;
                ClubEdge ce;
                ce.w_idx = e0 + (pf - b); ce.s_idx = -1; ce.target_row = rf; ce.pad_ = 0;
                co.push_back((uint32_t)std::min<uint64_t>(con, 0xffffffffull));
                con += s->h_ncl[rf];
                ed.push_back(ce);
            }
            for (const int32_t* pf = b; pf != e; ++pf) {                 // :45-64
                const int32_t *b2, *e2;
                if (!s->adj_of(*pf, &b2, &e2)) continue;
                if (s->row_of(*pf) < 0) continue;
                const int64_t f0 = b2 - s->adj.data();
                for (const int32_t* p = b2; p != e2; ++p) {
                    if (*p == uid) continue;
                    const int rc = s->row_of(*p);
                    if (rc < 0) continue;                                // no profile
                    ClubEdge ce;
                    ce.w_idx = e0 + (pf - b); ce.s_idx = f0 + (p - b2); ce.target_row = rc; ce.pad_ = 0;
                    co.push_back((uint32_t)std::min<uint64_t>(con, 0xffffffffull));
                    con += s->h_ncl[rc];
                    ed.push_back(ce);
                }
            }
            u_total[i] = con;
        });
        std::vector<ClubEdge> edges;
        std::vector<int32_t> ueb(nu + 1, 0), ucon(nu + 1, 0);
        std::vector<uint32_t> con_off;
        for (int i = 0; i < nu; ++i) {
            const uint64_t con = u_total[i];
            if (con > 0x7fffffffull) return fail(PK_ERR_UNSUPPORTED, "club contribution stream too large");
            edges.insert(edges.end(), u_edges[i].begin(), u_edges[i].end());
            con_off.insert(con_off.end(), u_con[i].begin(), u_con[i].end());
            std::vector<ClubEdge>().swap(u_edges[i]);
            std::vector<uint32_t>().swap(u_con[i]);
            if (edges.size() > 0x7fffffffull) return fail(PK_ERR_UNSUPPORTED, "club batch too large (split the batch)");
            ueb[i + 1] = (int32_t)edges.size();
            const int64_t capu = (int64_t)1 << log2_table(std::max<uint64_t>(con, 1));
            if ((int64_t)ucon[i] + capu > 0x7fffffffll) return fail(PK_ERR_UNSUPPORTED, "club batch too large (split the batch)");
            ucon[i + 1] = ucon[i] + (int32_t)capu;
        }
        const int64_t total_slots = ucon[nu];
        if (!edges.empty()) {
            TRY(s->d_edges.reserve(edges.size()));
            CK(cudaMemcpyAsync(s->d_edges.p, edges.data(), edges.size() * sizeof(ClubEdge), cudaMemcpyHostToDevice, s->stream));
            TRY(upload(s, s->d_u32a, con_off));
            TRY(upload(s, s->d_i32a, ueb));
            TRY(upload(s, s->d_i32b, ucon));
            TRY(upload(s, s->d_i32c, urow));
            TRY(s->d_i32d.reserve((size_t)total_slots));
            TRY(s->d_keys_a.reserve((size_t)total_slots));
            TRY(s->d_f64a.reserve((size_t)total_slots));
            TRY(s->d_keys_b.reserve((size_t)nu * K));
            CK(cudaMemsetAsync(s->d_keys_a.p, 0, (size_t)total_slots * 8, s->stream));
            CK(cudaMemsetAsync(s->d_f64a.p, 0, (size_t)total_slots * 8, s->stream));
            ClubParams C;
            C.rec = s->d_rec.p; C.rec_off = s->d_rec_off.p; C.fas = s->d_edge_fas.p; C.edges = s->d_edges.p;
            C.con_off = s->d_u32a.p; C.user_edge_begin = s->d_i32a.p; C.user_slot_begin = s->d_i32b.p;
            C.user_row = s->d_i32c.p; C.con_slot = s->d_i32d.p; C.slot_key = s->d_keys_a.p; C.slot_acc = s->d_f64a.p;
            TRY(s->d_keys_c.reserve((size_t)total_slots)); TRY(s->d_i32e.reserve((size_t)nu));
            C.out_keys = s->d_keys_c.p; C.out_cnt = s->d_i32e.p;
            clubs_scatter_kernel<<<nu, kClubThreads, 0, s->stream>>>(C);
            ++s->stat_launches;
            CK(cudaGetLastError());
            TRY(launch_slice_topk(s, s->d_keys_c.p, s->d_i32b.p, nu, K, s->d_keys_b.p, s->d_i32e.p));
            std::vector<uint64_t> uk((size_t)nu * K);
            CK(cudaMemcpyAsync(uk.data(), s->d_keys_b.p, uk.size() * 8, cudaMemcpyDeviceToHost, s->stream));
            CK(cudaStreamSynchronize(s->stream));
            for (int i = 0; i < nu; ++i) memcpy(&keys[(size_t)uq[i] * K], &uk[(size_t)i * K], (size_t)K * 8);
        }
    }
    memcpy(h_keys, keys.data(), (size_t)nq * K * 8);
    return PK_OK;
}

int recommend_clubs(pk_store* s, int nq, const int32_t* q_uid, int k, uint64_t* h_keys) {
    // the cache mirrors the LOADED adjacency: rows replaced behind the store (pk_set_adj_row) bypass it
    if (s->edge_cache_valid && s->adj_override.empty()) return recommend_clubs_cached(s, nq, q_uid, k, h_keys);
    const int K = std::max(k, 1);
    std::vector<CollabUser> users;
    Plan plan;
    std::vector<WorkItem> items;
    std::vector<int32_t> cand_rows;
    int64_t rows = 0;
    TRY(collab_prepare(s, nq, q_uid, 0, false, users, plan, items, cand_rows, &rows));
    s->stat_groups = (int64_t)plan.groups.size();
    std::vector<uint64_t> keys((size_t)std::max(nq, 1) * K, 0ull);
    const int nu = (int)users.size();
    if (nu > 0 && rows > 0) {
        TRY(upload_plan(s, plan, false));
        TRY(upload(s, s->d_cand_rows, cand_rows));
        TRY(s->d_f32a.reserve((size_t)rows * 32));
        ScanParams P = base_scan_params(s, K);
        P.cand_rows = s->d_cand_rows.p;
        P.out_f = s->d_f32a.p;
        TRY((run_scan<EPI_STORE, false>(s, plan, P, items)));

        // edge stream per user, in the reference's visiting order: 1-hop entries (target = friend), then
        // 2-hop entries (target = fof); each entry carries the indices of the FAS values it needs.
        std::vector<ClubEdge> edges;
        std::vector<int32_t> ueb(nu + 1, 0), ucon(nu + 1, 0), urow(nu);
        std::vector<uint32_t> con_off;   // per edge: first contribution index (relative to the user)
        // phase 1 (host threads): every user's edge stream
        std::vector<std::vector<ClubEdge>> u_edges((size_t)nu);
        std::vector<std::vector<uint32_t>> u_con((size_t)nu);
        std::vector<uint64_t> u_total((size_t)nu, 0);
        parallel_for(nu, [&](int i) {
            const CollabUser& u = users[i];
            std::vector<ClubEdge>& ed = u_edges[i];
            std::vector<uint32_t>& co = u_con[i];
            std::unordered_map<int32_t, int32_t> cand_idx;
            cand_idx.reserve(u.cands.size() * 2);
            for (size_t c = 0; c < u.cands.size(); ++c) cand_idx.emplace(u.cands[c], (int32_t)c);
            const int64_t ncand = (int64_t)u.cands.size();
            uint64_t con = 0;
            auto w_index = [&](int slot) { return ((int64_t)u.w_out0 + slot) * 32; };
            for (size_t fi = 0; fi < u.friends.size(); ++fi) {          // src/recommender_clubs.cpp:34-43
                const int sl = u.fr_slot[fi];
                if (sl < 0) continue;
                ClubEdge e;
                e.w_idx = w_index(sl); e.s_idx = -1; e.target_row = u.slot_row[sl];
                co.push_back((uint32_t)con);
                con += s->h_ncl[e.target_row];
                ed.push_back(e);
            }
            for (size_t fi = 0; fi < u.friends.size(); ++fi) {          // :45-64
                const int sl = u.fr_slot[fi];
                const int32_t *b2, *e2;
                if (!s->adj_of(u.friends[fi], &b2, &e2)) continue;
                if (sl < 0) continue;
                for (const int32_t* p = b2; p != e2; ++p) {
                    if (*p == u.uid) continue;
                    auto ic = cand_idx.find(*p);
                    if (ic == cand_idx.end()) continue;                 // no profile
                    ClubEdge e;
                    e.w_idx = w_index(sl);
                    e.s_idx = ((int64_t)u.mat_out0 + (int64_t)(sl >> 5) * ncand + ic->second) * 32 + (sl & 31);
                    e.target_row = u.cand_row[ic->second];
                    co.push_back((uint32_t)std::min<uint64_t>(con, 0xffffffffull));
                    con += s->h_ncl[e.target_row];
                    ed.push_back(e);
                }
            }
            u_total[i] = con;
        });
        // phase 2: concatenate in user order
        for (int i = 0; i < nu; ++i) {
            urow[i] = users[i].row;
            const uint64_t con = u_total[i];
            if (con > 0x7fffffffull) return fail(PK_ERR_UNSUPPORTED, "club contribution stream too large");
            edges.insert(edges.end(), u_edges[i].begin(), u_edges[i].end());
            con_off.insert(con_off.end(), u_con[i].begin(), u_con[i].end());
            std::vector<ClubEdge>().swap(u_edges[i]);
            std::vector<uint32_t>().swap(u_con[i]);
            if (edges.size() > 0x7fffffffull) return fail(PK_ERR_UNSUPPORTED, "club batch too large (split the batch)");
            ueb[i + 1] = (int32_t)edges.size();
            // table capacity for the user's distinct clubs (power of two >= 2 * contributions)
            const int64_t capu = (int64_t)1 << log2_table(std::max<uint64_t>(con, 1));
            if ((int64_t)ucon[i] + capu > 0x7fffffffll) return fail(PK_ERR_UNSUPPORTED, "club batch too large (split the batch)");
            ucon[i + 1] = ucon[i] + (int32_t)capu;
        }
        const int64_t total_slots = ucon[nu];
        if (!edges.empty()) {
            TRY(s->d_edges.reserve(edges.size()));
            CK(cudaMemcpyAsync(s->d_edges.p, edges.data(), edges.size() * sizeof(ClubEdge), cudaMemcpyHostToDevice, s->stream));
            TRY(upload(s, s->d_u32a, con_off));
            TRY(upload(s, s->d_i32a, ueb));
            TRY(upload(s, s->d_i32b, ucon));
            TRY(upload(s, s->d_i32c, urow));
            // per contribution slot index: sized by the largest per-user contribution count is not known
            // per user on the device, so it shares the table-slot budget (contributions <= capacity / 2)
            TRY(s->d_i32d.reserve((size_t)total_slots));
            TRY(s->d_keys_a.reserve((size_t)total_slots));          // club id + 1 per slot (as uint64: low 32 bits), then keys
            TRY(s->d_f64a.reserve((size_t)total_slots));
            TRY(s->d_keys_b.reserve((size_t)nu * K));
            CK(cudaMemsetAsync(s->d_keys_a.p, 0, (size_t)total_slots * 8, s->stream));
            CK(cudaMemsetAsync(s->d_f64a.p, 0, (size_t)total_slots * 8, s->stream));
            ClubParams C;
            C.rec = s->d_rec.p; C.rec_off = s->d_rec_off.p; C.fas = s->d_f32a.p; C.edges = s->d_edges.p;
            C.con_off = s->d_u32a.p; C.user_edge_begin = s->d_i32a.p; C.user_slot_begin = s->d_i32b.p;
            C.user_row = s->d_i32c.p; C.con_slot = s->d_i32d.p; C.slot_key = s->d_keys_a.p; C.slot_acc = s->d_f64a.p;
            TRY(s->d_keys_c.reserve((size_t)total_slots)); TRY(s->d_i32e.reserve((size_t)nu));
            C.out_keys = s->d_keys_c.p; C.out_cnt = s->d_i32e.p;
            clubs_scatter_kernel<<<nu, kClubThreads, 0, s->stream>>>(C);
            ++s->stat_launches;
            CK(cudaGetLastError());
            TRY(launch_slice_topk(s, s->d_keys_c.p, s->d_i32b.p, nu, K, s->d_keys_b.p, s->d_i32e.p));
            std::vector<uint64_t> uk((size_t)nu * K);
            CK(cudaMemcpyAsync(uk.data(), s->d_keys_b.p, uk.size() * 8, cudaMemcpyDeviceToHost, s->stream));
            CK(cudaStreamSynchronize(s->stream));
            for (int i = 0; i < nu; ++i) memcpy(&keys[(size_t)users[i].q * K], &uk[(size_t)i * K], (size_t)K * 8);
        }
    }
    memcpy(h_keys, keys.data(), (size_t)nq * K * 8);
    return PK_OK;
}

// ---- pairs --------------------------------------------------------------------------------------------
// one lane per pair (pk_pairs.cuh): a_row / b_row on the device, scores (and per-field floats) written in pair order
template <bool FIELDS>
int pairs_launch(pk_store* s, int64_t n, const int32_t* d_a_row, const int32_t* d_b_row, float* d_out_f, float* d_out_fields,
                 const int32_t* d_key_id, uint64_t* d_out_keys, bool timed) {
    if (n <= 0) return PK_OK;
    PairParams P;
    memset(&P, 0, sizeof(P));
    P.rec = s->d_rec.p; P.rec_off = s->d_rec_off.p; P.norm = s->d_norm.p; P.tab_comp = s->d_tab_comp.p; P.tab_age = s->d_tab_age.p;
    P.a_row = d_a_row; P.b_row = d_b_row; P.n = n; P.T = s->T; P.nf = 3 + s->T + 2;
    P.out_f = d_out_f; P.out_fields = d_out_fields; P.key_id = d_key_id; P.out_keys = d_out_keys;
    if (timed) CK(cudaEventRecord(s->ev0, s->stream));
    fas_pairs_kernel<FIELDS><<<(unsigned)((n + 127) / 128), 128, 0, s->stream>>>(P);
    if (timed) CK(cudaEventRecord(s->ev1, s->stream));
    CK(cudaGetLastError());
    ++s->stat_launches;
    if (timed) { ++s->stat_scan_launches; s->scan_time_pending = true; }
    return PK_OK;
}

// host pairs in, host results out: out_f[i] = FAS or -1 (a side without a profile); FIELDS: out_fields[i * nf + j]
template <bool FIELDS>
int pairs_run(pk_store* s, int64_t n, const int32_t* a_uid, const int32_t* b_uid, std::vector<float>& out_f,
              std::vector<float>& out_fields) {
    s->stat_scan_ns = 0; s->stat_scan_launches = 0; s->stat_pairs = 0;
    const int nf = 3 + s->T + 2;
    out_f.assign((size_t)n, -1.0f);
    if (FIELDS) out_fields.assign((size_t)n * nf, -1.0f);
    const int64_t kBatch = FIELDS ? (1 << 18) : (1 << 22);      // bounded device buffers
    std::vector<int32_t> ra, rb;
    for (int64_t p0 = 0; p0 < n; p0 += kBatch) {
        const int64_t m = std::min(kBatch, n - p0);
        ra.resize((size_t)m); rb.resize((size_t)m);
        std::atomic<int64_t> valid{0};
        parallel_for((int)((m + 4095) / 4096), [&](int blk) {
            int64_t v = 0;
            for (int64_t i = (int64_t)blk * 4096, e = std::min(m, (int64_t)(blk + 1) * 4096); i < e; ++i) {
                ra[(size_t)i] = s->row_of(a_uid[p0 + i]);
                rb[(size_t)i] = s->row_of(b_uid[p0 + i]);
                v += (ra[(size_t)i] >= 0 && rb[(size_t)i] >= 0) ? 1 : 0;
            }
            valid += v;
        });
        s->stat_pairs += valid.load();
        TRY(upload(s, s->d_i32a, ra));
        TRY(upload(s, s->d_i32b, rb));
        TRY(s->d_f32a.reserve((size_t)m));
        if (FIELDS) TRY(s->d_fields.reserve((size_t)m * nf));
        TRY((pairs_launch<FIELDS>(s, m, s->d_i32a.p, s->d_i32b.p, s->d_f32a.p, FIELDS ? s->d_fields.p : nullptr, nullptr, nullptr)));
        CK(cudaMemcpyAsync(out_f.data() + p0, s->d_f32a.p, (size_t)m * 4, cudaMemcpyDeviceToHost, s->stream));
        if (FIELDS) CK(cudaMemcpyAsync(out_fields.data() + (size_t)p0 * nf, s->d_fields.p, (size_t)m * nf * 4, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        TRY(collect_scan_time(s));
    }
    return PK_OK;
}

}  // namespace

// =======================================================================================================
// C ABI
// =======================================================================================================
extern "C" {

const char* pk_last_error(void) { return g_err.c_str(); }
int pk_abi_version(void) { return PK_ABI_VERSION; }

int pk_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int pk_store_create(int device, pk_store** out) {
    if (!out) return fail(PK_ERR_ARG, "null out pointer");
    *out = nullptr;
    int ndev = pk_device_count();
    if (ndev <= 0) return fail(PK_ERR_CUDA, "no CUDA device: libpokec_b200 has no CPU path");
    if (device < 0 || device >= ndev) return fail(PK_ERR_ARG, "device %d out of range (%d visible)", device, ndev);
    CK(cudaSetDevice(device));
    pk_store* s = new pk_store();
    s->device = device;
    cudaError_t e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&s->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&s->ev1);
    if (e == cudaSuccess) e = cudaEventCreate(&s->ev2);
    if (e == cudaSuccess) e = cudaEventCreate(&s->ev3);
    for (int i = 0; i < 6 && e == cudaSuccess; ++i) e = cudaEventCreate(&s->pev[i]);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&s->max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    if (e != cudaSuccess) { delete s; return fail(PK_ERR_CUDA, "store init: %s", cudaGetErrorString(e)); }
    *out = s;
    return PK_OK;
}

void pk_store_destroy(pk_store* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    for (int r = 0; r < 7; ++r) if (s->peer_mapped[r]) cudaIpcCloseMemHandle(s->peer_mapped[r]);
    if (s->stream) cudaStreamSynchronize(s->stream);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->ev2) cudaEventDestroy(s->ev2);
    if (s->ev3) cudaEventDestroy(s->ev3);
    for (int i = 0; i < 6; ++i) if (s->pev[i]) cudaEventDestroy(s->pev[i]);
    cudaStream_t st = s->stream;
    delete s;
    if (st) cudaStreamDestroy(st);
}

int pk_store_load_world(pk_store* s, const pk_world* w) {
    if (!s || !w) return fail(PK_ERR_ARG, "null argument");
    const int n = w->n_users, T = w->n_text_cols, m = w->n_adj_rows;
    if (n < 0 || T < 0 || m < 0) return fail(PK_ERR_ARG, "negative size");
    if (T > kMaxT) return fail(PK_ERR_UNSUPPORTED, "%d text columns; at most %d are supported", T, kMaxT);
    if (n > 0 && (!w->uid || !w->pub || !w->comp || !w->gender || !w->age || !w->region || !w->clubs_ptr || !w->friends_ptr ||
                  !w->tok_ptr))
        return fail(PK_ERR_ARG, "null profile array");
    if (m > 0 && (!w->adj_uid || !w->adj_ptr)) return fail(PK_ERR_ARG, "null adjacency array");
    CK(cudaSetDevice(s->device));
    // every offset array starts at 0 and never decreases (a damaged world file or a caller bug must not turn into
    // out-of-bounds reads in the pack kernels or in the host's adjacency walk)
    auto monotone = [](const int64_t* p, int64_t cnt) {
        if (p[0] != 0) return false;
        for (int64_t i = 0; i < cnt; ++i) if (p[i + 1] < p[i]) return false;
        return true;
    };
    if (n > 0 && (!monotone(w->clubs_ptr, n) || !monotone(w->friends_ptr, n) || !monotone(w->tok_ptr, (int64_t)n * T)))
        return fail(PK_ERR_ARG, "clubs_ptr / friends_ptr / tok_ptr must start at 0 and be non-decreasing");
    if (m > 0 && !monotone(w->adj_ptr, m)) return fail(PK_ERR_ARG, "adj_ptr must start at 0 and be non-decreasing");
    const int64_t ncl = n ? w->clubs_ptr[n] : 0, nfr = n ? w->friends_ptr[n] : 0, nnz = n ? w->tok_ptr[(int64_t)n * T] : 0;
    if ((ncl > 0 && !w->clubs) || (nfr > 0 && !w->friends) || (nnz > 0 && (!w->tok_id || !w->tok_tf)) ||
        (m > 0 && w->adj_ptr[m] > 0 && !w->adj))
        return fail(PK_ERR_ARG, "null element array");
    // cells sorted by token id (the reference's cells are hash maps: order-free); ids must fit 26 bits
    std::vector<int32_t> tid, ttf;
    const int32_t *tok_id = w->tok_id, *tok_tf = w->tok_tf;
    bool sorted = true, tf_negative = false;
    int64_t max_cell = 0;
    for (int64_t c = 0; c < (int64_t)n * T; ++c) {
        const int64_t b = w->tok_ptr[c], e = w->tok_ptr[c + 1];
        if (e < b) return fail(PK_ERR_ARG, "tok_ptr not monotone");
        max_cell = std::max(max_cell, e - b);
        for (int64_t j = b; j < e; ++j) {
            if ((uint32_t)tok_id[j] >= (1u << kTokBits)) return fail(PK_ERR_UNSUPPORTED, "token id %d outside [0, 2^26)", tok_id[j]);
            if (j > b && tok_id[j] <= tok_id[j - 1]) sorted = false;
            if (w->tok_tf[j] < 0) tf_negative = true;
        }
    }
    if (!sorted) {
        tid.assign(tok_id, tok_id + nnz);
        ttf.assign(tok_tf, tok_tf + nnz);
        std::vector<std::pair<int32_t, int32_t>> tmp;
        for (int64_t c = 0; c < (int64_t)n * T; ++c) {
            const int64_t b = w->tok_ptr[c], e = w->tok_ptr[c + 1];
            if (e - b < 2 || std::is_sorted(tid.begin() + b, tid.begin() + e)) continue;
            tmp.clear();
            for (int64_t j = b; j < e; ++j) tmp.emplace_back(tid[j], ttf[j]);
            std::sort(tmp.begin(), tmp.end());
            for (int64_t j = b; j < e; ++j) { tid[j] = tmp[j - b].first; ttf[j] = tmp[j - b].second; }
        }
        tok_id = tid.data(); tok_tf = ttf.data();
    }
    s->n = n; s->T = T; s->nnz = nnz;
    s->tf_negative = tf_negative;
    s->max_cell = max_cell;
    s->wide_ids = false;
    for (int64_t j = 0, e = ncl; j < e && !s->wide_ids; ++j) s->wide_ids = (w->clubs[j] >> 31) != 0;
    for (int64_t j = 0, e = nfr; j < e && !s->wide_ids; ++j) s->wide_ids = (w->friends[j] >> 31) != 0;
    s->uid.assign(w->uid, w->uid + n);
    s->uid_sorted.resize(n);
    for (int i = 0; i < n; ++i) s->uid_sorted[i] = {w->uid[i], i};
    std::sort(s->uid_sorted.begin(), s->uid_sorted.end());
    for (int i = 1; i < n; ++i)
        if (s->uid_sorted[i].first == s->uid_sorted[i - 1].first) return fail(PK_ERR_ARG, "duplicate profile uid %d", s->uid_sorted[i].first);
    s->h_ncl.resize(n); s->h_nfr.resize(n); s->h_ntok.assign(n, 0);
    s->h_pub.assign(w->pub, w->pub + n); s->h_gender.assign(w->gender, w->gender + n);
    s->h_comp.assign(w->comp, w->comp + n); s->h_age.assign(w->age, w->age + n);
    s->max_comp = 0; s->max_age = 0;
    for (int i = 0; i < n; ++i) {
        s->h_ncl[i] = (uint32_t)(w->clubs_ptr[i + 1] - w->clubs_ptr[i]);
        s->h_nfr[i] = (uint32_t)(w->friends_ptr[i + 1] - w->friends_ptr[i]);
        s->max_comp = std::max(s->max_comp, w->comp[i]);
        s->max_age = std::max(s->max_age, w->age[i]);
    }
    s->adj_uid.assign(w->adj_uid, w->adj_uid + m);
    s->adj_ptr.assign(w->adj_ptr, w->adj_ptr + (m ? m + 1 : 0));
    if (m == 0) s->adj_ptr.assign(1, 0);
    s->adj.assign(w->adj, w->adj + (m ? w->adj_ptr[m] : 0));
    s->adj_row.clear();
    s->adj_row.reserve((size_t)m * 2);
    for (int r = 0; r < m; ++r) s->adj_row[w->adj_uid[r]] = r;
    s->adj_override.clear();
    s->graph_dirty = true;
    s->max_deg = 0;
    for (int r = 0; r < m; ++r) s->max_deg = std::max<int64_t>(s->max_deg, w->adj_ptr[r + 1] - w->adj_ptr[r]);
    s->row_direct.clear(); s->adj_direct.clear();
    {
        int64_t lo = INT64_MAX, hi = INT64_MIN;
        for (int i = 0; i < n; ++i) { lo = std::min<int64_t>(lo, w->uid[i]); hi = std::max<int64_t>(hi, w->uid[i]); }
        for (int r = 0; r < m; ++r) { lo = std::min<int64_t>(lo, w->adj_uid[r]); hi = std::max<int64_t>(hi, w->adj_uid[r]); }
        if (n + m > 0 && hi - lo < 8 * (int64_t)(n + m) + 4096) {
            s->uid_lo = (int32_t)lo;
            s->row_direct.assign((size_t)(hi - lo + 1), -1);
            s->adj_direct.assign((size_t)(hi - lo + 1), -1);
            for (int i = 0; i < n; ++i) s->row_direct[(size_t)(w->uid[i] - lo)] = i;
            for (int r = 0; r < m; ++r) s->adj_direct[(size_t)(w->adj_uid[r] - lo)] = r;
        }
    }

    TRY(upload(s, s->d_uid, w->uid, n)); TRY(upload(s, s->d_pub, w->pub, n)); TRY(upload(s, s->d_comp, w->comp, n));
    TRY(upload(s, s->d_gender, w->gender, n)); TRY(upload(s, s->d_age, w->age, n)); TRY(upload(s, s->d_region, w->region, (size_t)3 * n));
    std::vector<int64_t> one(1, 0);
    TRY(upload(s, s->d_clubs_ptr, n ? w->clubs_ptr : one.data(), (size_t)n + 1));
    TRY(upload(s, s->d_friends_ptr, n ? w->friends_ptr : one.data(), (size_t)n + 1));
    TRY(upload(s, s->d_tok_ptr, n ? w->tok_ptr : one.data(), (size_t)n * T + 1));
    {   // the packed records keep every user's club / friend ids ASCENDING (duplicates kept): the similarity only
        // counts members (src/recommender.cpp:119-128), and the lane-per-pair kernel merges the two lists
        std::vector<uint32_t> cs(w->clubs, w->clubs + ncl), fs(w->friends, w->friends + nfr);
        parallel_for((n + 1023) / 1024, [&](int blk) {
            for (int r = blk * 1024, e = std::min(n, (blk + 1) * 1024); r < e; ++r) {
                if (w->clubs_ptr[r + 1] - w->clubs_ptr[r] > 1) std::sort(cs.begin() + w->clubs_ptr[r], cs.begin() + w->clubs_ptr[r + 1]);
                if (w->friends_ptr[r + 1] - w->friends_ptr[r] > 1) std::sort(fs.begin() + w->friends_ptr[r], fs.begin() + w->friends_ptr[r + 1]);
            }
        });
        TRY(upload(s, s->d_clubs, cs.data(), (size_t)ncl)); TRY(upload(s, s->d_friends, fs.data(), (size_t)nfr));
        CK(cudaStreamSynchronize(s->stream));       // cs / fs are locals
    }
    TRY(upload(s, s->d_tok_id, tok_id, (size_t)nnz)); TRY(upload(s, s->d_tok_tf, tok_tf, (size_t)nnz));
    CK(cudaStreamSynchronize(s->stream));

    s->col_state.assign(T, 0);
    s->custom_tok.assign(T, {});
    s->custom_idf.assign(T, {});
    s->idf_dirty = true; s->store_dirty = true; s->edge_cache_valid = false;
    const int nf = kNumFixed + T;
    s->nmean.assign(nf, 0.f); s->nsd.assign(nf, 0.f); s->nhas.assign(nf, 0);
    if (w->norm_mean && w->norm_sd && w->norm_has)
        for (int f = 0; f < nf; ++f) { s->nmean[f] = w->norm_mean[f]; s->nsd[f] = w->norm_sd[f]; s->nhas[f] = w->norm_has[f]; }
    s->norm_dirty = true; s->edge_cache_valid = false;
    s->shard_begin = 0; s->shard_end = n;
    ++s->gen;
    s->job.valid = false;
    s->seed_b = s->seed_e = -1;
    s->loaded = true;
    return PK_OK;
}

int pk_store_set_option(pk_store* s, int which, int64_t value) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    switch (which) {
        case PK_OPT_PREFILTER_RESERVE_SMS:
            if (value < 0 || value > 64) return fail(PK_ERR_ARG, "reserve_sms out of range");
            s->reserve_sms = (int)value;
            return PK_OK;
        default: return fail(PK_ERR_ARG, "unknown option %d", which);
    }
}

int pk_store_set_shard(pk_store* s, int64_t row_begin, int64_t row_end) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    if (row_begin < 0 || row_end < row_begin || row_end > s->n) return fail(PK_ERR_ARG, "bad shard [%lld,%lld)", (long long)row_begin, (long long)row_end);
    s->shard_begin = row_begin; s->shard_end = row_end;
    ++s->gen;
    return PK_OK;
}

int pk_set_normalizers(pk_store* s, const float* mean, const float* sd, const uint8_t* has, int n_fields) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    if (n_fields != kNumFixed + s->T) return fail(PK_ERR_ARG, "n_fields = %d, expected %d", n_fields, kNumFixed + s->T);
    if (!mean || !sd) return fail(PK_ERR_ARG, "null normalizer array");
    for (int f = 0; f < n_fields; ++f) { s->nmean[f] = mean[f]; s->nsd[f] = sd[f]; s->nhas[f] = has ? has[f] : 1; }
    s->norm_dirty = true; s->edge_cache_valid = false;
    ++s->gen;
    return PK_OK;
}

int pk_compute_idf(pk_store* s, int double_log) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    for (int t = 0; t < s->T; ++t) { s->col_state[t] = 1; s->custom_tok[t].clear(); s->custom_idf[t].clear(); }
    s->double_log = double_log ? 1 : 0;
    s->idf_dirty = true; s->edge_cache_valid = false;
    ++s->gen;
    CK(cudaSetDevice(s->device));
    return rebuild_idf(s);
}

int pk_set_idf(pk_store* s, int col, const int32_t* token, const float* idf, int64_t n) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    if (col < 0 || col >= s->T) return fail(PK_ERR_ARG, "column %d out of range", col);
    if (n > 0 && (!token || !idf)) return fail(PK_ERR_ARG, "null idf arrays");
    if (n < 0) {
        s->col_state[col] = 0; s->custom_tok[col].clear(); s->custom_idf[col].clear();
    } else {
        for (int64_t i = 0; i < n; ++i)
            if ((uint32_t)token[i] >= (1u << kTokBits)) return fail(PK_ERR_UNSUPPORTED, "token id %d outside [0, 2^26)", token[i]);
        s->col_state[col] = 2;
        s->custom_tok[col].assign(token, token + n);
        s->custom_idf[col].assign(idf, idf + n);
    }
    s->idf_dirty = true; s->edge_cache_valid = false;
    ++s->gen;
    return PK_OK;
}

int pk_idf_size(pk_store* s, int col, int64_t* n) {
    if (!s || !n) return fail(PK_ERR_ARG, "null argument");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    if (col < 0 || col >= s->T) return fail(PK_ERR_ARG, "column %d out of range", col);
    if (s->col_state[col] == 0) { *n = -1; return PK_OK; }
    TRY(pk_get_idf(s, col, nullptr, nullptr, nullptr, 0));
    *n = s->last_idf_n;
    return PK_OK;
}

int pk_get_idf(pk_store* s, int col, int32_t* token, float* idf, int32_t* df, int64_t cap) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    if (col < 0 || col >= s->T) return fail(PK_ERR_ARG, "column %d out of range", col);
    CK(cudaSetDevice(s->device));
    TRY(rebuild_idf(s));
    s->last_idf_n = 0;
    if (s->col_state[col] == 0) return PK_OK;
    // count, then export
    DevBuf<unsigned long long> d_cnt;
    TRY(d_cnt.reserve(1));
    const size_t tcap = (size_t)1 << s->idf_log_cap;
    for (int pass = 0; pass < 2; ++pass) {
        const int64_t want = pass == 0 ? 0 : s->last_idf_n;
        TRY(s->d_i32a.reserve(std::max<int64_t>(want, 1)));
        TRY(s->d_i32b.reserve(std::max<int64_t>(want, 1)));
        TRY(s->d_f32a.reserve(std::max<int64_t>(want, 1)));
        CK(cudaMemsetAsync(d_cnt.p, 0, 8, s->stream));
        idf_export_kernel<<<(unsigned)((tcap + 255) / 256), 256, 0, s->stream>>>(idf_table(s), col, s->d_i32a.p, s->d_f32a.p, s->d_i32b.p,
                                                                               d_cnt.p, want);
        ++s->stat_launches;
        unsigned long long cnt = 0;
        CK(cudaMemcpyAsync(&cnt, d_cnt.p, 8, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        s->last_idf_n = (int64_t)cnt;
        if (pass == 0 && (cnt == 0 || !token)) return PK_OK;
    }
    const int64_t m = s->last_idf_n;
    std::vector<int32_t> ht(m), hd(m);
    std::vector<float> hv(m);
    CK(cudaMemcpyAsync(ht.data(), s->d_i32a.p, m * 4, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaMemcpyAsync(hd.data(), s->d_i32b.p, m * 4, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaMemcpyAsync(hv.data(), s->d_f32a.p, m * 4, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    std::vector<int64_t> ord(m);
    for (int64_t i = 0; i < m; ++i) ord[i] = i;
    std::sort(ord.begin(), ord.end(), [&](int64_t a, int64_t b) { return ht[a] < ht[b]; });
    for (int64_t i = 0; i < m && i < cap; ++i) {
        token[i] = ht[ord[i]];
        if (idf) idf[i] = hv[ord[i]];
        if (df) df[i] = hd[ord[i]];
    }
    return PK_OK;
}

int pk_set_adj_row(pk_store* s, int32_t uid, const int32_t* nb, int64_t n) {
    if (!s || n < 0 || (n > 0 && !nb)) return fail(PK_ERR_ARG, "bad argument");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    ++s->gen;          // a prepared full scan carries the exclusion ids of the old rows
    // a row equal to the loaded one needs no override (the drop-in re-sends the rows a query touches)
    auto it = s->adj_row.find(uid);
    if (it != s->adj_row.end()) {
        const int64_t b = s->adj_ptr[it->second], e = s->adj_ptr[it->second + 1];
        if (e - b == n && (n == 0 || memcmp(s->adj.data() + b, nb, (size_t)n * 4) == 0)) {
            s->adj_override.erase(uid);
            return PK_OK;
        }
    } else if (n == 0) {
        s->adj_override.erase(uid);   // no row and an empty row enumerate the same (no) candidates
        return PK_OK;
    }
    s->adj_override[uid].assign(nb, nb + n);
    return PK_OK;
}

int pk_pair_similarity(pk_store* s, int64_t n, const int32_t* a_uid, const int32_t* b_uid, float* out) {
    TRY(ensure_ready(s));
    if (n < 0 || (n > 0 && (!a_uid || !b_uid || !out))) return fail(PK_ERR_ARG, "bad argument");
    std::vector<float> f, fields;
    TRY(pairs_run<false>(s, n, a_uid, b_uid, f, fields));
    if (n > 0) memcpy(out, f.data(), (size_t)n * 4);
    return PK_OK;
}

int pk_pair_fields(pk_store* s, int64_t n, const int32_t* a_uid, const int32_t* b_uid, float* out) {
    TRY(ensure_ready(s));
    if (n < 0 || (n > 0 && (!a_uid || !b_uid || !out))) return fail(PK_ERR_ARG, "bad argument");
    std::vector<float> f, fields;
    TRY(pairs_run<true>(s, n, a_uid, b_uid, f, fields));
    const int nfo = 3 + s->T, nf = nfo + 2;
    for (int64_t i = 0; i < n; ++i)
        for (int j = 0; j < nfo; ++j) out[i * nfo + j] = fields[(size_t)i * nf + j];
    return PK_OK;
}

// compute_column_normalizers (src/utils.cpp:155-240) for a caller-supplied pair sample: per field the mean and
// the sample standard deviation of the RAW similarities (0 where a field is absent, raw-count cosine for every
// text column whatever idf tables are set), accumulated in pair order in double exactly as the reference does.
int pk_normalizers_from_pairs(pk_store* s, int64_t n, const int32_t* a_uid, const int32_t* b_uid, float* out_mean, float* out_sd) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    if (n < 0 || (n > 0 && (!a_uid || !b_uid)) || !out_mean || !out_sd) return fail(PK_ERR_ARG, "bad argument");
    CK(cudaSetDevice(s->device));
    const int T = s->T, nfo = 3 + T, nf = nfo + 2, NF = kNumFixed + T;
    for (int64_t i = 0; i < n; ++i)
        if (s->row_of(a_uid[i]) < 0 || s->row_of(b_uid[i]) < 0)
            return fail(PK_ERR_ARG, "pair %lld names a user without a profile", (long long)i);   // the reference's profiles.at() throws
    // raw-count weights: every column temporarily loses its idf table (src/recommender_similarity.cpp:103 path),
    // the packed store is rebuilt for the sample and again, lazily, for the next scoring call
    const std::vector<uint8_t> saved = s->col_state;
    ++s->gen;
    std::fill(s->col_state.begin(), s->col_state.end(), (uint8_t)0);
    s->idf_dirty = true; s->store_dirty = true;
    std::vector<float> vals((size_t)n * nfo);        // region, clubs, friends, text columns of every pair
    int rc = ensure_ready(s);
    if (rc == PK_OK) {
        std::vector<float> f, fields;
        rc = pairs_run<true>(s, n, a_uid, b_uid, f, fields);
        for (int64_t i = 0; rc == PK_OK && i < n; ++i) memcpy(&vals[(size_t)i * nfo], &fields[(size_t)i * nf], (size_t)nfo * 4);
    }
    s->col_state = saved;
    s->idf_dirty = true; s->store_dirty = true;
    if (rc != PK_OK) return rc;
    // per-field value of pair i, in the reference's field order (public, gender, completion, age, region, clubs,
    // friends, text columns); -1 from the kernel = field absent on a side = the reference's 0.0
    auto value = [&](int64_t i, int fld) -> double {
        const int ra = s->row_of(a_uid[i]), rb = s->row_of(b_uid[i]);
        switch (fld) {
            case 0: return (s->h_pub[ra] >= 0 && s->h_pub[rb] >= 0 && s->h_pub[ra] == s->h_pub[rb]) ? 1.0 : 0.0;
            case 1: return (s->h_gender[ra] >= 0 && s->h_gender[rb] >= 0 && s->h_gender[ra] == s->h_gender[rb]) ? 1.0 : 0.0;
            case 2: case 3: {
                const int x = fld == 2 ? s->h_comp[ra] : s->h_age[ra], y = fld == 2 ? s->h_comp[rb] : s->h_age[rb];
                if (x <= 0 || y <= 0) return 0.0;
                return (double)std::min(x, y) / (double)std::max(x, y);
            }
            default: {
                const float v = vals[(size_t)i * nfo + (fld - 4)];             // 4 region, 5 clubs, 6 friends, 7+t text
                return v < 0.0f ? 0.0 : (double)v;
            }
        }
    };
    parallel_for(NF, [&](int fld) {
        if (n == 0) { out_mean[fld] = 0.0f; out_sd[fld] = 1.0f; return; }
        double mu = 0.0;
        for (int64_t i = 0; i < n; ++i) mu += value(i, fld);
        mu /= (double)n;
        double sd = 1.0;
        if (n > 1) {
            double acc = 0.0;
            for (int64_t i = 0; i < n; ++i) { const double d = value(i, fld) - mu; acc += d * d; }
            sd = std::sqrt(acc / (double)(n - 1));
            if (sd == 0.0) sd = 1.0;
        }
        out_mean[fld] = (float)mu; out_sd[fld] = (float)sd;
    });
    return PK_OK;
}

int pk_edge_cache_build(pk_store* s) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    CK(cudaSetDevice(s->device));
    return edge_cache_build(s);
}

int pk_edge_cache_drop(pk_store* s) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    s->edge_cache_valid = false;
    return PK_OK;
}

int pk_edge_cache_row(pk_store* s, int32_t uid, float* out, int64_t cap, int64_t* n) {
    if (!s || !n || (cap > 0 && !out)) return fail(PK_ERR_ARG, "null argument");
    if (!s->edge_cache_valid) return fail(PK_ERR_STATE, "no valid edge cache (pk_edge_cache_build)");
    *n = 0;
    const int64_t i = (int64_t)uid - s->uid_lo;
    int32_t r = -1;
    if (!s->adj_direct.empty()) r = (i >= 0 && i < (int64_t)s->adj_direct.size()) ? s->adj_direct[(size_t)i] : -1;
    else { auto it = s->adj_row.find(uid); if (it != s->adj_row.end()) r = it->second; }
    if (r < 0) return PK_OK;
    const int64_t b = s->adj_ptr[r], e = s->adj_ptr[r + 1];
    *n = e - b;
    for (int64_t j = b; j < e && j - b < cap; ++j) out[j - b] = s->h_edge_fas[(size_t)j];
    return PK_OK;
}

int pk_pair_overlap(pk_store* s, int64_t n, const int32_t* a_uid, const int32_t* b_uid, int which, int32_t* out) {
    TRY(ensure_ready(s));
    if (n < 0 || which < 0 || which > 1 || (n > 0 && (!a_uid || !b_uid || !out))) return fail(PK_ERR_ARG, "bad argument");
    std::vector<float> f, fields;
    TRY(pairs_run<true>(s, n, a_uid, b_uid, f, fields));
    const int nf = 3 + s->T + 2;
    for (int64_t i = 0; i < n; ++i) {
        if (f[(size_t)i] < 0.0f) { out[i] = -1; continue; }          // a side without a profile
        float v = fields[(size_t)i * nf + 3 + s->T + which];
        int32_t iv;
        memcpy(&iv, &v, 4);
        // the slot keeps its -1.0f initialiser when the field is absent on either side: overlap 0
        out[i] = (v == -1.0f) ? 0 : iv;
    }
    return PK_OK;
}

int pk_fullscan_prepare(pk_store* s, int nq, const int32_t* q_uid, int k, int exclude_mode) {
    TRY(ensure_ready(s));
    TRY(check_topk_args(s, nq, q_uid, k));
    if (exclude_mode < PK_EXCLUDE_NONE || exclude_mode > PK_EXCLUDE_SELF_AND_ADJ) return fail(PK_ERR_ARG, "bad exclude_mode");
    return fullscan_prepare(s, nq, q_uid, k, exclude_mode);
}

int pk_fullscan_execute(pk_store* s) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    CK(cudaSetDevice(s->device));
    return fullscan_execute(s);
}

int pk_fullscan_execute_shortlists(pk_store* s) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    CK(cudaSetDevice(s->device));
    return fullscan_shortlists(s);
}

int pk_fullscan_theta_export(pk_store* s, void* handle64) {
    if (!s || !handle64) return fail(PK_ERR_ARG, "null argument");
    CK(cudaSetDevice(s->device));
    TRY(s->d_theta_x.reserve(2 * pk_store::kThetaCap));
    cudaIpcMemHandle_t h;
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    CK(cudaIpcGetMemHandle(&h, s->d_theta_x.p));
    memcpy(handle64, &h, 64);
    return PK_OK;
}

void* pk_fullscan_theta_ptr(pk_store* s) {
    if (!s) return nullptr;
    cudaSetDevice(s->device);
    if (s->d_theta_x.reserve(2 * pk_store::kThetaCap) != PK_OK) return nullptr;
    return s->d_theta_x.p;
}

int pk_fullscan_theta_peers(pk_store* s, const void* handles, int n_ranks, int rank, const void* const* d_ptrs) {
    if (!s || n_ranks < 1 || n_ranks > 8 || rank < 0 || rank >= n_ranks || (!handles && !d_ptrs)) return fail(PK_ERR_ARG, "bad argument");
    CK(cudaSetDevice(s->device));
    for (int r = 0; r < 7; ++r) {
        if (s->peer_mapped[r]) cudaIpcCloseMemHandle(s->peer_mapped[r]);
        s->peer_mapped[r] = nullptr; s->peer_theta[r] = nullptr;
    }
    s->n_peers = 0;
    s->exec_counter = 0;            // every rank restarts its execution counter here: the tags stay in step
    for (int r = 0; r < n_ranks; ++r) {
        if (r == rank) continue;
        void* p = nullptr;
        if (d_ptrs) p = const_cast<void*>(d_ptrs[r]);
        else {
            cudaIpcMemHandle_t h;
            memcpy(&h, (const char*)handles + (size_t)r * 64, 64);
            cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                for (int i = 0; i < s->n_peers; ++i) { if (s->peer_mapped[i]) cudaIpcCloseMemHandle(s->peer_mapped[i]); s->peer_mapped[i] = nullptr; s->peer_theta[i] = nullptr; }
                s->n_peers = 0;
                return fail(PK_ERR_CUDA, "cudaIpcOpenMemHandle(rank %d): %s (the scan runs without threshold sharing)", r, cudaGetErrorString(e));
            }
            s->peer_mapped[s->n_peers] = p;
        }
        if (!p) continue;
        s->peer_theta[s->n_peers++] = reinterpret_cast<const unsigned long long*>(p);
    }
    return PK_OK;
}

int pk_fullscan_shortlist_keys(pk_store* s, uint64_t* d_short) {
    if (!s || !d_short) return fail(PK_ERR_ARG, "null argument");
    FullscanJob& J = s->job;
    TRY(job_ok(s));
    if (!J.lean || J.nq == 0) return PK_OK;
    CK(cudaSetDevice(s->device));
    CK(cudaMemcpyAsync(d_short, J.keys_short.p, (size_t)J.nq * J.Ks * 8, cudaMemcpyDeviceToDevice, s->stream));
    return PK_OK;
}

int pk_fullscan_execute_finish(pk_store* s, const uint64_t* d_all_short, int n_lists, int rank, int n_ranks) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    CK(cudaSetDevice(s->device));
    return fullscan_finish(s, d_all_short, n_lists, rank, n_ranks);
}

int pk_fullscan_fetch(pk_store* s, int32_t* out_id, float* out_score, int32_t* out_cnt) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    if (s->job.valid && s->job.nq > 0 && (!out_cnt || (s->job.k > 0 && (!out_id || !out_score)))) return fail(PK_ERR_ARG, "null output");
    CK(cudaSetDevice(s->device));
    return fullscan_fetch(s, out_id, out_score, out_cnt);
}

int pk_fullscan_keys(pk_store* s, uint64_t* d_keys) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    TRY(job_ok(s));
    if (s->job.nq == 0 || s->job.k == 0) return PK_OK;
    if (!d_keys) return fail(PK_ERR_ARG, "null d_keys");
    CK(cudaMemcpyAsync(d_keys, s->job.keys_b.p, (size_t)s->job.nq * s->job.k * 8, cudaMemcpyDeviceToDevice, s->stream));
    return PK_OK;
}

int pk_fullscan_topk(pk_store* s, int nq, const int32_t* q_uid, int k, int exclude_mode, int32_t* out_id, float* out_score,
                     int32_t* out_cnt) {
    if (nq > 0 && (!out_cnt || (k > 0 && (!out_id || !out_score)))) return fail(PK_ERR_ARG, "null output");
    TRY(pk_fullscan_prepare(s, nq, q_uid, k, exclude_mode));
    TRY(fullscan_execute(s));
    return fullscan_fetch(s, out_id, out_score, out_cnt);
}

int pk_fullscan_local_keys(pk_store* s, int nq, const int32_t* q_uid, int k, int exclude_mode, uint64_t* d_keys) {
    TRY(pk_fullscan_prepare(s, nq, q_uid, k, exclude_mode));
    TRY(fullscan_execute(s));
    return pk_fullscan_keys(s, d_keys);
}

int pk_merge_keys(pk_store* s, const uint64_t* d_keys, int n_lists, int nq, int k, int32_t* out_id, float* out_score,
                  int32_t* out_cnt) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    if (n_lists < 1 || nq < 0 || k < 0 || k > kMaxK) return fail(PK_ERR_ARG, "bad n_lists/nq/k");
    if (nq == 0) return PK_OK;
    const bool device_only = !out_id && !out_score && !out_cnt;     // merge on the device, fetch nothing (asynchronous)
    if (k == 0) { for (int q = 0; q < nq && out_cnt; ++q) out_cnt[q] = 0; return PK_OK; }
    if (!d_keys || (!device_only && (!out_id || !out_score || !out_cnt))) return fail(PK_ERR_ARG, "null argument");
    if (n_lists > 128 * 8) return fail(PK_ERR_UNSUPPORTED, "at most 1024 lists");
    CK(cudaSetDevice(s->device));
    TRY(s->d_keys_c.reserve((size_t)nq * k));
    MergeParams M;
    memset(&M, 0, sizeof(M));
    M.keys = d_keys; M.n_lists = n_lists; M.nq = nq; M.K = k; M.stride_k = k; M.out_keys = s->d_keys_c.p;
    TRY(launch_merge(s, M, nq));
    if (device_only) return PK_OK;
    std::vector<uint64_t> keys((size_t)nq * k);
    CK(cudaMemcpyAsync(keys.data(), s->d_keys_c.p, keys.size() * 8, cudaMemcpyDeviceToHost, s->stream));
    s->stat_d2h += (int64_t)keys.size() * 8;
    CK(cudaStreamSynchronize(s->stream));
    decode_keys(keys.data(), nq, k, out_id, out_score, out_cnt);
    return PK_OK;
}

int pk_recommend(pk_store* s, int mode, int nq, const int32_t* q_uid, int k, int limit, int32_t* out_id, float* out_score,
                 int32_t* out_cnt) {
    TRY(ensure_ready(s));
    TRY(check_topk_args(s, nq, q_uid, k, kMaxListK));
    if (nq > 0 && (!out_cnt || (k > 0 && (!out_id || !out_score)))) return fail(PK_ERR_ARG, "null output");
    if (mode != PK_MODE_GRAPH && mode != PK_MODE_COLLAB && mode != PK_MODE_CLUBS) return fail(PK_ERR_ARG, "bad mode %d", mode);
    if (nq == 0) return PK_OK;
    if (k == 0) { for (int q = 0; q < nq; ++q) out_cnt[q] = 0; return PK_OK; }
    s->stat_scan_ns = 0; s->stat_scan_launches = 0; s->stat_pairs = 0; s->stat_groups = 0;
    std::vector<uint64_t> keys((size_t)nq * k, 0ull);
    if (mode == PK_MODE_CLUBS && device_lists_usable(s, mode, nq, 1)) TRY(recommend_clubs_device(s, nq, q_uid, k, keys.data()));
    else if (mode != PK_MODE_CLUBS && device_lists_usable(s, mode, nq, limit)) TRY(recommend_lists_device(s, mode, nq, q_uid, k, limit, keys.data()));
    else if (mode == PK_MODE_GRAPH) TRY(recommend_graph(s, nq, q_uid, k, limit, keys.data()));
    else if (mode == PK_MODE_COLLAB) TRY(recommend_collab(s, nq, q_uid, k, limit, keys.data()));
    else TRY(recommend_clubs(s, nq, q_uid, k, keys.data()));
    decode_keys(keys.data(), nq, k, out_id, out_score, out_cnt);
    return PK_OK;
}

int pk_candidates(pk_store* s, int mode, int32_t uid, int limit, int32_t* out, int64_t cap, int64_t* n) {
    if (!s || !n || (cap > 0 && !out)) return fail(PK_ERR_ARG, "null argument");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    std::vector<int32_t> c;
    if (mode == PK_MODE_GRAPH) gather_candidates(s, uid, limit, c);
    else if (mode == PK_MODE_COLLAB) {
        const int32_t *b = nullptr, *e = nullptr;
        if (s->adj_of(uid, &b, &e)) collab_candidates(s, uid, b, e, limit, true, c);
    } else return fail(PK_ERR_ARG, "bad mode %d", mode);
    *n = (int64_t)c.size();
    for (int64_t i = 0; i < (int64_t)c.size() && i < cap; ++i) out[i] = c[i];
    return PK_OK;
}

int pk_candidates_batch(pk_store* s, int mode, int nq, const int32_t* q_uid, int limit, int32_t* out_ids, int32_t* out_cnt) {
    if (!s || nq < 0 || (nq > 0 && (!q_uid || !out_ids || !out_cnt))) return fail(PK_ERR_ARG, "bad argument");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    if (mode != PK_MODE_GRAPH && mode != PK_MODE_COLLAB) return fail(PK_ERR_ARG, "bad mode %d", mode);
    if (limit < 1 || limit > 16384 || s->max_deg > 32768 || !s->adj_override.empty() || s->adj.size() > 0x7fffff00ull)
        return fail(PK_ERR_UNSUPPORTED, "the device expansion takes limits in [1, 16384] on an adjacency without overridden rows");
    if (nq == 0) return PK_OK;
    TRY(ensure_ready(s));
    for (int q0 = 0; q0 < nq; q0 += 8192) {
        const int m = std::min(8192, nq - q0);
        TRY(expand_batch(s, mode, m, q_uid + q0, limit, true));
        CK(cudaMemcpyAsync(out_ids + (size_t)q0 * limit, s->g_cand.p, (size_t)m * limit * 4, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaMemcpyAsync(out_cnt + q0, s->g_cnt.p, (size_t)m * 4, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
    }
    return PK_OK;
}

// ---- pk_candidates_batch_noalloc: the raw device expansion on a STATIC scratch array ------------------------------------
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  expand_kernel needs a per-warp hash
// set, its epoch tags, a work-queue cursor and a [nq] row output; here they live in one file-scope device array instead
// of the store's growable buffers.  Every GPU that loads the library holds the array (about 4 MB) until exit.
//   layout (8-byte words): [kNaWarps][kNaTabCap] tables | kNaWarps epochs (uint32) | cursor (4 ints) | kNaMaxQ rows (int32)
// The epoch-tagged tables never need clearing between calls (a call continues the epochs the last one left, exactly as
// the store's own scratch does between launches); the cursor must start at zero and is cleared on every call.  Calls
// share the array: two calls that overlap on the GPU (different stores / streams of one device) race.
constexpr int kNaCtas = 32;                                   // CTAs of the no-alloc launch
constexpr int kNaWarps = kNaCtas * kExpWarps;                 // 256 warps
constexpr uint32_t kNaTabCap = 2048;                          // slots per warp: >= 2 (limit + 32)
constexpr int kNaMaxLimit = (int)(kNaTabCap / 2) - 32;        // 992
constexpr int kNaMaxQ = 8192;                                 // query users per call
constexpr size_t kNaOffEpoch = (size_t)kNaWarps * kNaTabCap;  // in 8-byte words
constexpr size_t kNaOffNext = kNaOffEpoch + kNaWarps / 2;
constexpr size_t kNaOffRow = kNaOffNext + 32;
constexpr size_t kNaWords = kNaOffRow + kNaMaxQ / 2;
static __device__ __align__(256) unsigned long long pk_noalloc_scratch[kNaWords];

// In host code `pk_noalloc_scratch` is a host-side stand-in, not a device address: the device address is asked for
// once per device, with that device current, and cached.
static int noalloc_scratch_ptr(int device, unsigned long long** out) {
    constexpr int kMaxDev = 64;
    static unsigned long long* ptrs[kMaxDev];
    static std::atomic<int> have[kMaxDev];
    if (device < 0 || device >= kMaxDev) return fail(PK_ERR_UNSUPPORTED, "device ordinal %d outside [0, %d)", device, kMaxDev);
    if (!have[device].load(std::memory_order_acquire)) {
        CK(cudaSetDevice(device));
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, pk_noalloc_scratch));
        if (!p) return fail(PK_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        ptrs[device] = (unsigned long long*)p;                 // two racing callers store the same value
        have[device].store(1, std::memory_order_release);
    }
    *out = ptrs[device];
    return PK_OK;
}

int pk_candidates_batch_noalloc(pk_store* s, int mode, int nq, const int32_t* d_q_uid, int limit, int32_t* d_out_ids,
                                int32_t* d_out_cnt) {
    if (!s || nq < 0 || (nq > 0 && (!d_q_uid || !d_out_ids || !d_out_cnt))) return fail(PK_ERR_ARG, "bad argument");
    if (!s->loaded) return fail(PK_ERR_STATE, "no world loaded");
    if (mode != PK_MODE_GRAPH && mode != PK_MODE_COLLAB) return fail(PK_ERR_ARG, "bad mode %d", mode);
    if (limit < 1 || limit > kNaMaxLimit)
        return fail(PK_ERR_UNSUPPORTED, "pk_candidates_batch_noalloc takes limits in [1, %d] (static scratch); got %d", kNaMaxLimit, limit);
    if (nq > kNaMaxQ)
        return fail(PK_ERR_UNSUPPORTED, "pk_candidates_batch_noalloc takes at most %d query users per call; got %d", kNaMaxQ, nq);
    if (!s->adj_override.empty() || s->adj.size() > 0x7fffff00ull)
        return fail(PK_ERR_UNSUPPORTED, "the device expansion needs an adjacency without overridden rows");
    if (s->graph_dirty)
        return fail(PK_ERR_STATE, "the adjacency is not resident on the device yet and this entry point allocates nothing: "
                                  "call pk_candidates_batch (or a batched pk_recommend) once first");
    if (nq == 0) return PK_OK;
    CK(cudaSetDevice(s->device));
    unsigned long long* scratch = nullptr;
    TRY(noalloc_scratch_ptr(s->device, &scratch));
    int* next = reinterpret_cast<int*>(scratch + kNaOffNext);
    CK(cudaMemsetAsync(next, 0, 16, s->stream));
    ExpandParams X;
    memset(&X, 0, sizeof(X));
    X.G = dev_graph(s);
    X.q_uid = d_q_uid; X.nq = nq; X.mode = mode; X.limit = limit; X.raw = 1; X.slots_only = 0;
    X.scratch = scratch; X.epoch = reinterpret_cast<uint32_t*>(scratch + kNaOffEpoch); X.tab_cap = kNaTabCap;
    X.next_query = next; X.cand_cap = limit;
    X.q_row = reinterpret_cast<int32_t*>(scratch + kNaOffRow); X.cand_out = d_out_ids; X.cand_cnt = d_out_cnt;
    const int ctas = std::max(1, std::min(kNaCtas, (nq + kExpWarps - 1) / kExpWarps));
    expand_kernel<<<ctas, kExpWarps * 32, 0, s->stream>>>(X);
    ++s->stat_launches;
    CK(cudaGetLastError());
    return PK_OK;
}

int pk_debug_prefilter(pk_store* s, int nq, const int32_t* q_uid, int64_t row_begin, int64_t row_end, float* out_approx,
                       float* out_exact) {
    TRY(ensure_ready(s));
    if (nq < 1 || nq > 32 || !q_uid || !out_approx || !out_exact || row_begin < 0 || row_end > s->n || row_end < row_begin)
        return fail(PK_ERR_ARG, "bad argument");
    TRY(rebuild_lean(s));
    Plan plan;
    GroupDesc g = empty_group();
    for (int q = 0; q < nq; ++q) {
        const int r = s->row_of(q_uid[q]);
        if (r < 0) return fail(PK_ERR_ARG, "uid %d has no profile", q_uid[q]);
        g.row[q] = r; g.qidx[q] = q;
    }
    g.n_lanes = nq;
    finish_group(s, plan, g, nullptr, 0, true, 0);
    if (!plan.groups[0].lean_in_smem || !plan.groups[0].in_smem) return fail(PK_ERR_UNSUPPORTED, "group tables exceed shared memory");
    const int64_t rows = row_end - row_begin;
    if (rows == 0) return PK_OK;
    TRY(upload_plan(s, plan, false));
    WorkItem it;
    it.group = 0; it.cand_begin = (int32_t)row_begin; it.cand_end = (int32_t)row_end; it.out_slot = 0;
    std::vector<WorkItem> items(1, it);
    TRY(s->d_f32a.reserve((size_t)rows * 32));
    std::vector<float> tmp((size_t)rows * 32);
    for (int pass = 0; pass < 2; ++pass) {
        if (pass == 0) {
            TRY(upload(s, s->d_items, items));
            PreParams Q = base_pre_params(s);
            Q.groups = s->d_groups.p; Q.items = s->d_items.p; Q.arena = s->d_arena.p;
            Q.smem_arena_bytes = plan.max_smem_lean; Q.out_f = s->d_f32a.p;
            TRY(launch_prefilter(s, Q, 1, plan.max_smem_lean, 0, true));
        } else {
            ScanParams P = base_scan_params(s, 1);
            P.out_f = s->d_f32a.p;
            TRY((run_scan<EPI_STORE, false>(s, plan, P, items)));
        }
        CK(cudaMemcpyAsync(tmp.data(), s->d_f32a.p, tmp.size() * 4, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
        float* out = pass == 0 ? out_approx : out_exact;
        for (int64_t r = 0; r < rows; ++r)
            for (int q = 0; q < nq; ++q) out[r * nq + q] = tmp[(size_t)r * 32 + q];
    }
    return PK_OK;
}

int pk_sync(pk_store* s) {
    if (!s) return fail(PK_ERR_ARG, "null store");
    CK(cudaSetDevice(s->device));
    CK(cudaStreamSynchronize(s->stream));
    TRY(collect_scan_time(s));
    return PK_OK;
}

void* pk_stream(pk_store* s) { return s ? (void*)s->stream : nullptr; }

int64_t pk_stat(pk_store* s, int which) {
    if (!s) return -1;
    switch (which) {
        case PK_STAT_KERNEL_LAUNCHES: return s->stat_launches;
        case PK_STAT_PAIRS_SCORED: return s->stat_pairs;
        case PK_STAT_STORE_BYTES: return (int64_t)s->store_u4 * 16;
        case PK_STAT_SCAN_KERNEL_NS: return s->stat_scan_ns;
        case PK_STAT_SCAN_LAUNCHES: return s->stat_scan_launches;
        case PK_STAT_GROUPS: return s->stat_groups;
        case PK_STAT_SHARD_RECORD_BYTES:
            if (s->h_rec_off.size() < (size_t)s->n + 1) return 0;
            return (int64_t)(s->h_rec_off[s->shard_end] - s->h_rec_off[s->shard_begin]) * 16;
        case PK_STAT_H2D_BYTES: return s->stat_h2d;
        case PK_STAT_D2H_BYTES: return s->stat_d2h;
        case PK_STAT_PREFILTER: return s->job.valid && s->job.lean ? 1 : 0;
        case PK_STAT_FALLBACK_QUERIES: return s->stat_fallback;
        case PK_STAT_RESCORE_NS: return s->stat_rescore_ns;
        case PK_STAT_SHARD_LEAN_BYTES: {
            if (s->lean_dirty) return 0;
            int64_t b = 0;
            for (int64_t r = s->shard_begin; r < s->shard_end; ++r) b += 16 * (int64_t)lean_record_u4(s->h_ncl[r], s->h_nfr[r], s->h_ntok[r]);
            return b;
        }
        case PK_STAT_SHORTLIST: return s->job.valid && s->job.lean ? s->job.Ks : 0;
        case PK_STAT_MARGIN_E9: return (int64_t)((double)s->margin * 1e9);
        case PK_STAT_EDGE_CACHE: return s->edge_cache_valid ? (int64_t)s->h_edge_fas.size() : 0;
        case PK_STAT_PREP_NS: return s->stat_prep_ns;
        case PK_STAT_MERGE_NS: return s->stat_merge_ns;
        default: return -1;
    }
}

// ---- binary profile store on disk (include/pokec_b200.h) ------------------------------------------------
// One file = header + section table + the 19 arrays of a pk_world, each 64-byte aligned, native little-endian.  The
// reader maps the file read-only and hands out a pk_world whose pointers point into the mapping: no parse, no copy
// before pk_store_load_world streams the arrays to the device.
namespace {
constexpr uint64_t kWorldMagic = 0x31444c52574b5000ull;   // "\0PKWRLD1"
constexpr int kWorldSections = 19;
struct WorldFileHeader {
    uint64_t magic;
    uint32_t version, endian_tag;          // 1, 0x01020304
    int32_t n_users, n_text_cols, n_adj_rows, n_sections;
    uint64_t file_bytes;
    uint64_t off[kWorldSections], bytes[kWorldSections];
    uint64_t payload_sum;                  // hash of the section contents (payload_hash), verified by pk_world_open
    uint64_t header_sum;                   // FNV-1a of everything above
};
uint64_t fnv1a(const void* p, size_t n) {
    const unsigned char* b = (const unsigned char*)p;
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}
struct Section { const void* p; uint64_t bytes; };
// content hash of a section: four interleaved multiply-xor lanes over 8-byte words (memory speed, ~1 s for the
// 3 GB of a 1.63 M-user world), the tail bytes folded in last
uint64_t hash_bytes(const void* p, uint64_t n) {
    const unsigned char* b = (const unsigned char*)p;
    uint64_t h[4] = {0x9E3779B97F4A7C15ull, 0xC2B2AE3D27D4EB4Full, 0x165667B19E3779F9ull, 0x27D4EB2F165667C5ull};
    const uint64_t words = n / 8;
    uint64_t i = 0;
    for (; i + 4 <= words; i += 4)
        for (int l = 0; l < 4; ++l) {
            uint64_t w;
            memcpy(&w, b + (i + l) * 8, 8);
            h[l] = (h[l] ^ w) * 0x100000001B3ull;
            h[l] ^= h[l] >> 29;
        }
    uint64_t r = h[0];
    for (int l = 1; l < 4; ++l) r = (r ^ h[l]) * 0xFF51AFD7ED558CCDull;
    for (uint64_t j = i * 8; j < n; ++j) { r ^= b[j]; r *= 0x100000001B3ull; }
    return r ^ (r >> 32) ^ n;
}
uint64_t payload_hash(const Section* sec) {
    uint64_t r = 0xCBF29CE484222325ull;
    for (int i = 0; i < kWorldSections; ++i) r = (r ^ (sec[i].bytes ? hash_bytes(sec[i].p, sec[i].bytes) : (uint64_t)i)) * 0x100000001B3ull;
    return r;
}
// element counts follow from the header ints and the last entries of the three offset arrays
bool world_sections(const pk_world* w, Section* sec) {
    const int64_t n = w->n_users, T = w->n_text_cols, m = w->n_adj_rows, nf = PK_NUM_FIXED + T;
    if (n < 0 || T < 0 || m < 0) return false;
    if (n > 0 && (!w->clubs_ptr || !w->friends_ptr || !w->tok_ptr)) return false;
    if (m > 0 && !w->adj_ptr) return false;
    const int64_t ncl = n ? w->clubs_ptr[n] : 0, nfr = n ? w->friends_ptr[n] : 0, ntk = n ? w->tok_ptr[n * T] : 0,
                  nadj = m ? w->adj_ptr[m] : 0;
    if (ncl < 0 || nfr < 0 || ntk < 0 || nadj < 0) return false;
    const Section v[kWorldSections] = {
        {w->uid, (uint64_t)n * 4}, {w->pub, (uint64_t)n * 4}, {w->comp, (uint64_t)n * 4}, {w->gender, (uint64_t)n * 4},
        {w->age, (uint64_t)n * 4}, {w->region, (uint64_t)n * 12},
        {w->clubs_ptr, n ? (uint64_t)(n + 1) * 8 : 0}, {w->clubs, (uint64_t)ncl * 4},
        {w->friends_ptr, n ? (uint64_t)(n + 1) * 8 : 0}, {w->friends, (uint64_t)nfr * 4},
        {w->tok_ptr, n ? (uint64_t)(n * T + 1) * 8 : 0}, {w->tok_id, (uint64_t)ntk * 4}, {w->tok_tf, (uint64_t)ntk * 4},
        {w->adj_uid, (uint64_t)m * 4}, {w->adj_ptr, m ? (uint64_t)(m + 1) * 8 : 0}, {w->adj, (uint64_t)nadj * 4},
        {w->norm_mean, (uint64_t)nf * 4}, {w->norm_sd, (uint64_t)nf * 4}, {w->norm_has, (uint64_t)nf}};
    for (int i = 0; i < kWorldSections; ++i) {
        if (v[i].bytes && !v[i].p) return false;
        sec[i] = v[i];
    }
    return true;
}
}  // namespace

struct pk_world_file {
    void* map = nullptr;
    size_t bytes = 0;
};

int pk_world_save(const pk_world* w, const char* path) {
    if (!w || !path) return fail(PK_ERR_ARG, "pk_world_save: null argument");
    Section sec[kWorldSections];
    if (!world_sections(w, sec)) return fail(PK_ERR_ARG, "pk_world_save: inconsistent world (null array or negative size)");
    WorldFileHeader H;
    memset(&H, 0, sizeof(H));
    H.magic = kWorldMagic; H.version = 2; H.endian_tag = 0x01020304u;
    H.n_users = w->n_users; H.n_text_cols = w->n_text_cols; H.n_adj_rows = w->n_adj_rows; H.n_sections = kWorldSections;
    uint64_t off = (sizeof(H) + 63) & ~63ull;
    for (int i = 0; i < kWorldSections; ++i) { H.off[i] = off; H.bytes[i] = sec[i].bytes; off = (off + sec[i].bytes + 63) & ~63ull; }
    H.file_bytes = off;
    H.payload_sum = payload_hash(sec);
    H.header_sum = fnv1a(&H, offsetof(WorldFileHeader, header_sum));
    const std::string tmp = std::string(path) + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return fail(PK_ERR_ARG, "pk_world_save: cannot create %s", tmp.c_str());
    bool ok = fwrite(&H, sizeof(H), 1, f) == 1;
    uint64_t pos = sizeof(H);
    static const char zeros[64] = {0};
    for (int i = 0; ok && i < kWorldSections; ++i) {
        if (pos < H.off[i]) { ok = fwrite(zeros, 1, H.off[i] - pos, f) == H.off[i] - pos; pos = H.off[i]; }
        if (ok && sec[i].bytes) { ok = fwrite(sec[i].p, 1, sec[i].bytes, f) == sec[i].bytes; pos += sec[i].bytes; }
    }
    if (ok && pos < H.file_bytes) ok = fwrite(zeros, 1, H.file_bytes - pos, f) == H.file_bytes - pos;
    ok = (fclose(f) == 0) && ok;
    if (!ok || rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return fail(PK_ERR_ARG, "pk_world_save: write to %s failed", path); }
    return PK_OK;
}

int pk_world_open(const char* path, pk_world_file** out, pk_world* view) {
    if (!path || !out || !view) return fail(PK_ERR_ARG, "pk_world_open: null argument");
    *out = nullptr;
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return fail(PK_ERR_ARG, "pk_world_open: cannot open %s", path);
    struct stat st;
    if (fstat(fd, &st) != 0 || (size_t)st.st_size < sizeof(WorldFileHeader)) { close(fd); return fail(PK_ERR_ARG, "pk_world_open: %s is not a world file (too short)", path); }
    void* map = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (map == MAP_FAILED) return fail(PK_ERR_ARG, "pk_world_open: mmap of %s failed", path);
    const WorldFileHeader& H = *reinterpret_cast<const WorldFileHeader*>(map);
    const char* why = nullptr;
    if (H.magic != kWorldMagic) why = "bad magic";
    else if (H.version != 2 || H.n_sections != kWorldSections) why = "unsupported version";
    else if (H.endian_tag != 0x01020304u) why = "written on a machine of the other byte order";
    else if (H.header_sum != fnv1a(&H, offsetof(WorldFileHeader, header_sum))) why = "header checksum mismatch";
    else if (H.file_bytes != (uint64_t)st.st_size) why = "truncated (size differs from the header)";
    else if (H.n_users < 0 || H.n_text_cols < 0 || H.n_adj_rows < 0) why = "negative count";
    if (!why)
        for (int i = 0; i < kWorldSections; ++i)
            if ((H.off[i] & 63) || H.off[i] > H.file_bytes || H.bytes[i] > H.file_bytes - H.off[i]) { why = "section outside the file"; break; }
    pk_world w;
    memset(&w, 0, sizeof(w));
    if (!why) {
        const char* base = reinterpret_cast<const char*>(map);
        auto at = [&](int i) -> const void* { return H.bytes[i] ? base + H.off[i] : nullptr; };
        w.n_users = H.n_users; w.n_text_cols = H.n_text_cols; w.n_adj_rows = H.n_adj_rows;
        w.uid = (const int32_t*)at(0); w.pub = (const int32_t*)at(1); w.comp = (const int32_t*)at(2);
        w.gender = (const int32_t*)at(3); w.age = (const int32_t*)at(4); w.region = (const int32_t*)at(5);
        w.clubs_ptr = (const int64_t*)at(6); w.clubs = (const uint32_t*)at(7);
        w.friends_ptr = (const int64_t*)at(8); w.friends = (const uint32_t*)at(9);
        w.tok_ptr = (const int64_t*)at(10); w.tok_id = (const int32_t*)at(11); w.tok_tf = (const int32_t*)at(12);
        w.adj_uid = (const int32_t*)at(13); w.adj_ptr = (const int64_t*)at(14); w.adj = (const int32_t*)at(15);
        w.norm_mean = (const float*)at(16); w.norm_sd = (const float*)at(17); w.norm_has = (const uint8_t*)at(18);
        // the section sizes must be the ones the counts and offset arrays imply
        const int64_t n = H.n_users, T = H.n_text_cols, m = H.n_adj_rows;
        const uint64_t fixed[6] = {(uint64_t)n * 4, (uint64_t)n * 4, (uint64_t)n * 4, (uint64_t)n * 4, (uint64_t)n * 4, (uint64_t)n * 12};
        for (int i = 0; i < 6 && !why; ++i) if (H.bytes[i] != fixed[i]) why = "scalar section size";
        if (!why && (H.bytes[6] != (n ? (uint64_t)(n + 1) * 8 : 0) || H.bytes[8] != H.bytes[6] ||
                     H.bytes[10] != (n ? (uint64_t)(n * T + 1) * 8 : 0) || H.bytes[14] != (m ? (uint64_t)(m + 1) * 8 : 0) ||
                     H.bytes[13] != (uint64_t)m * 4)) why = "offset section size";
        if (!why) {
            Section sec[kWorldSections];
            if (!world_sections(&w, sec)) why = "inconsistent arrays";
            else for (int i = 0; i < kWorldSections; ++i) if (sec[i].bytes != H.bytes[i]) { why = "array size differs from its offsets"; break; }
            // the contents themselves (POKEC_B200_WORLD_VERIFY=0 skips the pass for callers that trust the file)
            const char* ev = getenv("POKEC_B200_WORLD_VERIFY");
            if (!why && !(ev && ev[0] == '0') && payload_hash(sec) != H.payload_sum) why = "payload checksum mismatch (damaged contents)";
        }
    }
    if (why) { munmap(map, (size_t)st.st_size); return fail(PK_ERR_ARG, "pk_world_open: %s: %s", path, why); }
    pk_world_file* f = new pk_world_file;
    f->map = map; f->bytes = (size_t)st.st_size;
    *out = f;
    *view = w;
    return PK_OK;
}

void pk_world_close(pk_world_file* f) {
    if (!f) return;
    if (f->map) munmap(f->map, f->bytes);
    delete f;
}


}  // extern "C"
