// pk_build.cuh -- kernels that build the device-resident profile store:
//   df histogram (replaces the counting loop of Recommender::compute_idf_from_profiles,
//   reference src/recommender.cpp:43-66, and of TFIDFIndex::build, src/tfidf_index.cpp:7-30),
//   record sizing and packing (precomputes the tf*idf weights and L2 norms that
//   tfidf_cosine_for_column, src/recommender.cpp:68-117, recomputes on every call).
#pragma once
#include "pk_layout.h"

namespace pk {

struct RawWorld {
    int n, T;
    const int32_t *uid, *pub, *comp, *gender, *age, *region;
    const int64_t* clubs_ptr;
    const uint32_t* clubs;
    const int64_t* friends_ptr;
    const uint32_t* friends;
    const int64_t* tok_ptr;
    const int32_t* tok_id;
    const int32_t* tok_tf;
};

// open-addressing table (col << 26 | token) -> {df, idf}
struct IdfTable {
    uint32_t* key;   // 0xFFFFFFFF = empty
    uint32_t* df;
    float* idf;
    int log_cap;
};
constexpr uint32_t kIdfEmpty = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t hash_u32(uint32_t key, int log_size) {
    return (key * 0x9E3779B1u) >> (32 - log_size);
}

// returns the slot of `key`, inserting it if needed; -1 when the table is full
__device__ __forceinline__ int64_t idf_insert(const IdfTable& I, uint32_t k) {
    const uint32_t mask = (1u << I.log_cap) - 1u;
    uint32_t slot = hash_u32(k, I.log_cap);
    for (uint32_t probe = 0; probe <= mask; ++probe) {
        uint32_t cur = I.key[slot];
        if (cur == k) return slot;
        if (cur == kIdfEmpty) {
            uint32_t old = atomicCAS(&I.key[slot], kIdfEmpty, k);
            if (old == kIdfEmpty || old == k) return slot;
        }
        slot = (slot + 1) & mask;
    }
    return -1;
}

__device__ __forceinline__ int64_t idf_find(const IdfTable& I, uint32_t k) {
    const uint32_t mask = (1u << I.log_cap) - 1u;
    uint32_t slot = hash_u32(k, I.log_cap);
    for (uint32_t probe = 0; probe <= mask; ++probe) {
        uint32_t cur = I.key[slot];
        if (cur == k) return slot;
        if (cur == kIdfEmpty) return -1;
        slot = (slot + 1) & mask;
    }
    return -1;
}

// one thread per (user, column) cell: df[token] += 1 for every token of the cell
// (a cell is a map, so each token is counted once per profile: src/recommender.cpp:50-57)
__global__ void df_count_kernel(RawWorld W, IdfTable I, const uint8_t* __restrict__ col_computed, int* overflow) {
    const int64_t cell = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n_cells = (int64_t)W.n * W.T;
    if (cell >= n_cells) return;
    const int t = (int)(cell % W.T);
    if (!col_computed[t]) return;
    const int64_t b = W.tok_ptr[cell], e = W.tok_ptr[cell + 1];
    for (int64_t j = b; j < e; ++j) {
        uint32_t k = ((uint32_t)t << kTokBits) | (uint32_t)W.tok_id[j];
        int64_t slot = idf_insert(I, k);
        if (slot < 0) { atomicExch(overflow, 1); return; }
        atomicAdd(&I.df[slot], 1u);
    }
}

// idf of every occupied slot from the host-evaluated table idf_of_df[df]
__global__ void idf_fill_kernel(IdfTable I, const float* __restrict__ idf_of_df, int64_t max_df) {
    const int64_t slot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= (1ll << I.log_cap)) return;
    if (I.key[slot] == kIdfEmpty) return;
    int64_t d = I.df[slot];
    if (d > max_df) d = max_df;
    I.idf[slot] = idf_of_df[d];
}

// caller-supplied idf entries (Recommender::set_tfidf_index, src/recommender.cpp:35-37)
__global__ void idf_custom_kernel(IdfTable I, int col, const int32_t* __restrict__ token, const float* __restrict__ idf,
                                  int64_t n, int* overflow) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t k = ((uint32_t)col << kTokBits) | (uint32_t)token[i];
    int64_t slot = idf_insert(I, k);
    if (slot < 0) { atomicExch(overflow, 1); return; }
    I.idf[slot] = idf[i];
    I.df[slot] = 0;
}

// compaction of one column's entries for read-back (pk_get_idf)
__global__ void idf_export_kernel(IdfTable I, int col, int32_t* token, float* idf, int32_t* df, unsigned long long* count,
                                  int64_t cap) {
    const int64_t slot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= (1ll << I.log_cap)) return;
    uint32_t k = I.key[slot];
    if (k == kIdfEmpty || (int)(k >> kTokBits) != col) return;
    unsigned long long pos = atomicAdd(count, 1ull);
    if ((int64_t)pos < cap) {
        token[pos] = (int32_t)(k & ((1u << kTokBits) - 1u));
        idf[pos] = I.idf[slot];
        df[pos] = (int32_t)I.df[slot];
    }
}

// one warp per row: packed size (uint4 units) and token count
__global__ void record_size_kernel(RawWorld W, uint32_t* __restrict__ size_u4, uint32_t* __restrict__ n_tok_out) {
    const int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (gw >= W.n) return;
    const int64_t row = gw;
    const int T = W.T;
    const int64_t c0 = row * T;
    int ne = 0;
    for (int t = lane; t < T; t += 32) ne += (W.tok_ptr[c0 + t + 1] > W.tok_ptr[c0 + t]) ? 1 : 0;
    for (int o = 16; o; o >>= 1) ne += __shfl_xor_sync(0xffffffffu, ne, o);
    if (lane == 0) {
        uint32_t ncl = (uint32_t)(W.clubs_ptr[row + 1] - W.clubs_ptr[row]);
        uint32_t nfr = (uint32_t)(W.friends_ptr[row + 1] - W.friends_ptr[row]);
        uint32_t nt = (uint32_t)(W.tok_ptr[c0 + T] - W.tok_ptr[c0]);
        size_u4[row] = record_u4(ncl, nfr, (uint32_t)ne, nt);
        n_tok_out[row] = nt;
    }
}

// one warp per row: write the packed record (layout: pk_layout.h)
__global__ void pack_kernel(RawWorld W, IdfTable I, const uint8_t* __restrict__ col_has_idf,
                            const uint64_t* __restrict__ rec_off, uint4* __restrict__ rec) {
    const int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (gw >= W.n) return;
    const int64_t row = gw;
    const int T = W.T;
    const int64_t c0 = row * T;
    uint32_t* out = reinterpret_cast<uint32_t*>(rec + rec_off[row]);

    const int64_t cb = W.clubs_ptr[row], fb = W.friends_ptr[row], tb = W.tok_ptr[c0];
    const uint32_t ncl = (uint32_t)(W.clubs_ptr[row + 1] - cb);
    const uint32_t nfr = (uint32_t)(W.friends_ptr[row + 1] - fb);
    const uint32_t ntok = (uint32_t)(W.tok_ptr[c0 + T] - tb);
    uint32_t len[2];
    for (int h = 0; h < 2; ++h) {
        int t = lane + 32 * h;
        len[h] = (t < T) ? (uint32_t)(W.tok_ptr[c0 + t + 1] - W.tok_ptr[c0 + t]) : 0u;
    }
    const uint32_t m0 = __ballot_sync(0xffffffffu, len[0] > 0), m1 = __ballot_sync(0xffffffffu, len[1] > 0);
    const uint32_t ncells = __popc(m0) + __popc(m1);
    const uint32_t o_clubs = kHdrWords, o_friends = o_clubs + pad4(ncl), o_cells = o_friends + pad4(nfr),
                   o_keys = o_cells + 4 * ncells, o_w = o_keys + pad4(ntok);

    if (lane == 0) {
        const int32_t r0 = W.region[3 * row], r1 = W.region[3 * row + 1], r2 = W.region[3 * row + 2];
        out[0] = (uint32_t)W.uid[row];
        out[1] = (uint32_t)W.pub[row];
        out[2] = (uint32_t)W.gender[row];
        out[3] = (uint32_t)W.comp[row];
        out[4] = (uint32_t)W.age[row];
        out[5] = (uint32_t)r0; out[6] = (uint32_t)r1; out[7] = (uint32_t)r2;
        out[8] = ncl; out[9] = nfr; out[10] = ncells; out[11] = ntok;
        out[12] = m0; out[13] = m1;
        out[14] = (uint32_t)((r0 >= 0) + (r1 >= 0) + (r2 >= 0));
        out[15] = 0u;
        reinterpret_cast<double*>(out + 16)[0] = sqrt((double)ncl);
        reinterpret_cast<double*>(out + 16)[1] = sqrt((double)nfr);
    }
    for (uint32_t i = lane; i < pad4(ncl); i += 32) out[o_clubs + i] = (i < ncl) ? W.clubs[cb + i] : 0u;
    for (uint32_t i = lane; i < pad4(nfr); i += 32) out[o_friends + i] = (i < nfr) ? W.friends[fb + i] : 0u;
    for (uint32_t i = ntok + lane; i < pad4(ntok); i += 32) out[o_keys + i] = 0u;
    if (lane == 0 && (ntok & 1u)) reinterpret_cast<double*>(out + o_w)[ntok] = 0.0;

    const uint32_t lt = (1u << lane) - 1u;
    for (int h = 0; h < 2; ++h) {
        if (len[h] == 0) continue;
        const int t = lane + 32 * h;
        const uint32_t k = h ? (__popc(m0) + __popc(m1 & lt)) : __popc(m0 & lt);
        const int64_t b = W.tok_ptr[c0 + t];
        const uint32_t start = (uint32_t)(b - tb);
        const bool has_idf = col_has_idf[t] != 0;
        double acc = 0.0;
        for (uint32_t i = 0; i < len[h]; ++i) {
            const uint32_t id = (uint32_t)W.tok_id[b + i];
            const uint32_t key = ((uint32_t)t << kTokBits) | id;
            double w = (double)W.tok_tf[b + i];
            if (has_idf) {
                int64_t slot = idf_find(I, key);
                float idf = slot >= 0 ? I.idf[slot] : 1.0f;      // src/recommender.cpp:78
                w = __dmul_rn(w, (double)idf);
            }
            acc = __dadd_rn(acc, __dmul_rn(w, w));
            out[o_keys + start + i] = key;
            reinterpret_cast<double*>(out + o_w)[start + i] = w;
        }
        out[o_cells + 4 * k + 0] = (uint32_t)t;
        out[o_cells + 4 * k + 1] = start + len[h];
        reinterpret_cast<double*>(out + o_cells + 4 * k + 2)[0] = sqrt(acc);
    }
}

}  // namespace pk
