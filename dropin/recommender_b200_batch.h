// recommender_b200_batch.h -- SURVEY.md section 8 row f1: many users per call.
//
// The reference's `Recommender` answers one user per member call (include/recommender.h:27-35) and its server
// loop calls the four paths once per `USER` line (src/api_cli.cpp:200-252).  The device library scores whole
// batches (pk_recommend, include/pokec_b200.h); this header is the small reference-side surface that lets a
// caller hand over a batch without touching include/recommender.h: a free function next to the drop-in
// `Recommender` (dropin/recommender_b200.cpp) plus the mode names of the `BATCH` line that
// dropin/api_cli_batch.inc adds to api_cli's line loop.
#pragma once

#include "recommender.h"

#include <string>
#include <utility>
#include <vector>

namespace pokec_b200 {

// "graph" | "collaborative" | "interest" | "clubs" (the keys of api_cli's "recommendations" object,
// src/api_cli.cpp:213-236) -> PK_MODE_*; -1 for anything else.
int batch_mode_from_name(const std::string& name);

// out[i] is exactly what the single-user member returns for users[i] with the same topk / candidate_limit
// (recommend_graph_registration, recommend_collaborative, recommend_by_interest, recommend_clubs_collab):
// same ids in the same order, same float scores.  Unknown users and users without neighbours get an empty list.
std::vector<std::vector<std::pair<int, float>>> recommend_batch(const Recommender& rec, int pk_mode, const std::vector<int>& users,
                                                                int topk, int candidate_limit);

}  // namespace pokec_b200
