/* pokec_world.h -- flat view of the reference's in-memory model, shared by the CUDA library
 * (include/pokec_b200.h), the CPU oracle (oracle/) and the compiled-reference driver.
 *
 * It is the array form of
 *   std::unordered_map<int, UserProfile>          (reference include/user_profile.h:10-20)
 *   std::unordered_map<int, std::vector<int>>     (reference include/recommender.h:49, adj_list)
 *   field_normalizers / column_normalizers        (reference include/recommender.h:51-52)
 * All pointers are HOST pointers owned by the caller; callees copy what they keep.
 */
#ifndef POKEC_WORLD_H
#define POKEC_WORLD_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PK_NUM_FIXED 7 /* public, gender, completion, age, region, clubs, friends
                          (reference src/recommender_similarity.cpp:12,38-91) */

typedef struct pk_world {
    int32_t n_users;            /* profiles->size() */
    int32_t n_text_cols;        /* T = text_columns.size(); total_possible = 7 + T */
    const int32_t* uid;         /* [n] profile keys, unique */
    const int32_t* pub;         /* [n] public_flag, <0 = missing */
    const int32_t* comp;        /* [n] completion_percentage, <=0 = missing */
    const int32_t* gender;      /* [n] gender, <0 = missing */
    const int32_t* age;         /* [n] age, <=0 = missing */
    const int32_t* region;      /* [n*3] region_parts, <0 = missing */
    const int64_t* clubs_ptr;   /* [n+1] */
    const uint32_t* clubs;      /* UserProfile::clubs, original order (duplicates allowed) */
    const int64_t* friends_ptr; /* [n+1] */
    const uint32_t* friends;    /* UserProfile::friends, original order */
    const int64_t* tok_ptr;     /* [n*T+1] cells row-major (user, column) */
    const int32_t* tok_id;      /* token ids, unique inside a cell (they are map keys) */
    const int32_t* tok_tf;      /* term counts */
    int32_t n_adj_rows;         /* adj_list->size() */
    const int32_t* adj_uid;     /* [m] adjacency row keys, unique */
    const int64_t* adj_ptr;     /* [m+1] */
    const int32_t* adj;         /* neighbour uids, original order */
    const float* norm_mean;     /* [7+T] normalizer mean per field (7 fixed, then columns) */
    const float* norm_sd;       /* [7+T] normalizer sd */
    const uint8_t* norm_has;    /* [7+T] 1 = the map has an entry for this key */
} pk_world;

/* recommendation modes (reference include/recommender.h:24-28) */
enum {
    PK_MODE_GRAPH = 0,   /* recommend_graph_registration == recommend_by_interest */
    PK_MODE_COLLAB = 1,  /* recommend_collaborative */
    PK_MODE_CLUBS = 2    /* recommend_clubs_collab */
};

/* exclusion rule of the full scan (the extension of SURVEY.md section 0 fact 2) */
enum {
    PK_EXCLUDE_NONE = 0,
    PK_EXCLUDE_SELF = 1,
    PK_EXCLUDE_SELF_AND_ADJ = 2 /* reference src/recommender_graph.cpp:46-52 */
};

#ifdef __cplusplus
}
#endif
#endif
