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
#include "user_profile.h"

#include <cstdint>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

namespace pokec_b200 {

// "graph" | "collaborative" | "interest" | "clubs" (the keys of api_cli's "recommendations" object,
// src/api_cli.cpp:213-236) -> PK_MODE_*; "fullscan" -> kBatchFullscan; -1 for anything else.
// kBatchFullscan is the path BASELINE.json's north_star names: FAS top-k of the user against EVERY profile instead of
// the friends-of-friends candidate set, minus the user and the existing friends (pk_fullscan_topk; topk <= kBatchMaxK).
constexpr int kBatchFullscan = 100;
constexpr int kBatchMaxK = 256;
int batch_mode_from_name(const std::string& name);

// out[i] is exactly what the single-user member returns for users[i] with the same topk / candidate_limit
// (recommend_graph_registration, recommend_collaborative, recommend_by_interest, recommend_clubs_collab):
// same ids in the same order, same float scores.  Unknown users and users without neighbours get an empty list.
std::vector<std::vector<std::pair<int, float>>> recommend_batch(const Recommender& rec, int pk_mode, const std::vector<int>& users,
                                                                int topk, int candidate_limit);

// ---- the reference's containers <-> one binary world file -------------------------------
// (pk_world_save / pk_world_open, include/pokec_b200.h).  Replaces the CSV round trip of save_users_encoded /
// load_users_encoded (src/user_loader.cpp:10-96, capped at 100 000 users) and adjacency.csv
// (src/graph_builder.cpp:39-75) for a server that restarts on unchanged data.
//   save_world_file  writes rec's profiles (rows in uid order), adjacency and normalizers; columns are the list given
//                    to rec.set_text_columns.  false (message on stderr) when the file cannot be written.
//   load_world_file  clears and fills the caller's maps; returns the number of text columns, -1 on a missing /
//                    damaged file (message on stderr).  The optional vectors receive the [7 + T] normalizer arrays
//                    (fixed fields in the order public, gender, completion, age, region, clubs, friends; then columns).
bool save_world_file(const Recommender& rec, const std::string& path);
int load_world_file(const std::string& path, std::unordered_map<int, UserProfile>& profiles,
                    std::unordered_map<int, std::vector<int>>& adj_list, std::vector<float>* norm_mean = nullptr,
                    std::vector<float>* norm_sd = nullptr, std::vector<uint8_t>* norm_has = nullptr);

}  // namespace pokec_b200
