// recommender_internal_sim.h -- the one definition of the friend struct that the reference header declares
// (`friend struct ::RecommenderInternalSim`, include/recommender.h:15,70) and no reference source defines.
// The drop-in uses it to read the private column list from free functions (recommender_b200_batch.h); the test
// driver oracle/ref_driver.cpp, when it is compiled INTO the drop-in library (-DPK_DROPIN_BUILD), includes this
// same header for its per-field accessors, so that the library holds a single definition of the class.
#pragma once

#include "recommender.h"

#include <array>
#include <string>
#include <unordered_map>
#include <vector>

struct RecommenderInternalSim {
    static const std::vector<std::string>& columns(const Recommender& r) { return r.text_columns_internal; }
    // private per-field helpers (include/recommender.h:59-66), each a separate float rounding point for the tests
    static float set_sim(const std::vector<uint32_t>& a, const std::vector<uint32_t>& b) { return Recommender::vec_set_similarity(a, b); }
    static float region_sim(const std::array<int, 3>& a, const std::array<int, 3>& b) { return Recommender::region_similarity_local(a, b); }
    static float raw_cos(const std::unordered_map<int, int>& a, const std::unordered_map<int, int>& b) {
        return Recommender::cosine_counts_maps_local(a, b);
    }
    static float tfidf_cos(const Recommender& r, const std::unordered_map<int, int>& a, const std::unordered_map<int, int>& b,
                           const std::unordered_map<int, float>& idf) {
        return r.tfidf_cosine_for_column(a, b, idf);
    }
};
