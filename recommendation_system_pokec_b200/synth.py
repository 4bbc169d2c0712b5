"""Seeded synthetic "Pokec-shaped" world.

Facts taken from the reference: the 49 text column names (config/text_columns.txt), uids are
positive ints (src/user_loader.cpp:44 skips 0), edges are directed and kept in file order
(src/graph_builder.cpp:8-21), UserProfile::friends is the adjacency row (src/encoder.cpp:117-124),
region has 3 hierarchical parts (src/vocab_builder.cpp:100-121), ages of 0 are filled with the
median (src/user_loader.cpp:131-140).  Dataset scale from BASELINE.json: 1.63 M users, ~30.6 M
directed edges (mean out-degree 18.75), Zipfian vocabulary per text column.  Everything else
(fill rates, vocabulary sizes, club counts, degree tail) is an assumption, parameterised here and
recorded in `World.meta` so runs are reproducible.
"""
from __future__ import annotations

import numpy as np

from .world import World, NUM_FIXED

TEXT_COLUMNS = (
    "I_am_working_in_field spoken_languages hobbies I_most_enjoy_good_food pets body_type "
    "my_eyesight eye_color hair_color hair_type completed_level_of_education favourite_color "
    "relation_to_smoking relation_to_alcohol sign_in_zodiac on_pokec_i_am_looking_for love_is_for_me "
    "relation_to_casual_sex my_partner_should_be marital_status children relation_to_children "
    "I_like_movies I_like_watching_movie I_like_music I_mostly_like_listening_to_music "
    "the_idea_of_good_evening I_like_specialties_from_kitchen fun I_am_going_to_concerts "
    "my_active_sports my_passive_sports profession I_like_books life_style music cars politics "
    "relationships art_culture hobbies_interests science_technologies computers_internet education "
    "sport movies travelling health companies_brands").split()

# categorical columns: small closed vocabularies, (mostly) one token per user
CATEGORICAL = {
    "I_am_working_in_field", "body_type", "my_eyesight", "eye_color", "hair_color", "hair_type",
    "completed_level_of_education", "favourite_color", "relation_to_smoking", "relation_to_alcohol",
    "sign_in_zodiac", "love_is_for_me", "relation_to_casual_sex", "marital_status", "children",
    "relation_to_children", "I_like_watching_movie", "I_mostly_like_listening_to_music",
    "I_am_going_to_concerts", "life_style", "pets", "spoken_languages",
}

POKEC_N = 1_632_803
POKEC_MEAN_DEG = 18.75
SEED = 0x504F4B4543  # "POKEC"


def _zipf_sampler(V, s):
    w = np.arange(1, V + 1, dtype=np.float64) ** (-s)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    return cdf


def _csr_from_rows(row, val, n):
    """rows (already grouped in the wanted order) -> ptr"""
    cnt = np.bincount(row, minlength=n).astype(np.int64)
    ptr = np.zeros(n + 1, np.int64)
    np.cumsum(cnt, out=ptr[1:])
    return ptr, val


def _dedup_keep_order(row, val):
    """drop duplicate (row,val) pairs keeping first occurrences and the original order."""
    key = row.astype(np.int64) * (1 << 32) + val.astype(np.int64)
    _, first = np.unique(key, return_index=True)
    first.sort()
    return row[first], val[first]


def make_world(n=100_000, seed=SEED, T=None, mean_deg=POKEC_MEAN_DEG, max_deg=None,
               uid_gap=False, text_scale=1.0) -> World:
    """Generate a world of `n` users.  `T` < 49 truncates the column list (small tests).
    `uid_gap=True` makes uids non-dense (tests the uid->row map)."""
    rng = np.random.default_rng(seed)
    cols = list(TEXT_COLUMNS if T is None else TEXT_COLUMNS[:T])
    T = len(cols)
    if max_deg is None:
        max_deg = int(min(8000, max(8, n // 4)))

    if uid_gap:
        uid = np.sort(rng.choice(np.arange(1, 3 * n + 1), size=n, replace=False)).astype(np.int32)
    else:
        uid = np.arange(1, n + 1, dtype=np.int32)

    # ---- scalars --------------------------------------------------------------
    pub = (rng.random(n) < 0.6).astype(np.int32)
    pub[rng.random(n) < 0.01] = -1
    gender = (rng.random(n) < 0.5).astype(np.int32)
    gender[rng.random(n) < 0.01] = -1
    comp = np.clip(np.rint(rng.beta(2.0, 2.5, n) * 100), 0, 100).astype(np.int32)
    comp[rng.random(n) < 0.01] = 0
    age = np.clip(np.rint(rng.gamma(6.0, 4.0, n) + 8), 9, 112).astype(np.int32)
    missing_age = rng.random(n) < 0.25
    real = age[~missing_age]
    med = 0
    if real.size:
        srt = np.sort(real)
        m = srt.size
        med = int(srt[m // 2]) if m % 2 else int((int(srt[m // 2 - 1]) + int(srt[m // 2])) // 2)
    age[missing_age] = med  # reference: fill_missing_ages (user_loader.cpp:131-140)

    r1 = rng.integers(0, 10, n)
    r2 = r1 * 20 + rng.integers(0, 20, n)
    r3 = r2 * 3 + rng.integers(0, 3, n)
    region = np.stack([r1, r2, r3], 1).astype(np.int32)
    region[rng.random((n, 3)) < 0.1] = -1

    # ---- directed graph --------------------------------------------------------
    sigma = 1.15
    mu = np.log(mean_deg * 1.10) - sigma * sigma / 2   # 1.10: compensates zero rows + dedup
    deg = np.minimum(np.rint(rng.lognormal(mu, sigma, n)), min(max_deg, n - 1)).astype(np.int64)
    deg[rng.random(n) < 0.04] = 0
    src = np.repeat(np.arange(n, dtype=np.int64), deg)
    E0 = src.shape[0]
    local = rng.random(E0) < 0.55
    # local: neighbours close in index space (communities -> shared friends);
    # global: popularity-weighted (Zipf) targets (hubs)
    off = np.rint(rng.laplace(0.0, 40.0, E0)).astype(np.int64)
    off[off == 0] = 1
    dst_local = (src + off) % n
    pop_cdf = _zipf_sampler(n, 0.8)
    perm = rng.permutation(n)
    dst_glob = perm[np.minimum(np.searchsorted(pop_cdf, rng.random(E0)), n - 1)]
    dst = np.where(local, dst_local, dst_glob)
    keep = dst != src
    src, dst = src[keep], dst[keep]
    src, dst = _dedup_keep_order(src, dst)
    order = np.argsort(src, kind="stable")          # group by source, keep (random) file order
    src, dst = src[order], dst[order]
    out_deg = np.bincount(src, minlength=n).astype(np.int64)
    has_row = out_deg > 0
    adj_uid = uid[has_row]
    adj_ptr = np.zeros(int(has_row.sum()) + 1, np.int64)
    np.cumsum(out_deg[has_row], out=adj_ptr[1:])
    adj = uid[dst].astype(np.int32)
    friends_ptr = np.zeros(n + 1, np.int64)
    np.cumsum(out_deg, out=friends_ptr[1:])
    friends = uid[dst].astype(np.uint32)

    # ---- clubs --------------------------------------------------------------------
    C = max(50, n // 8)
    ncl = np.where(rng.random(n) < 0.4, 0, np.minimum(1 + rng.geometric(1 / 7.5, n), 200))
    crow = np.repeat(np.arange(n, dtype=np.int64), ncl)
    club_cdf = _zipf_sampler(C, 1.0)
    cval = np.minimum(np.searchsorted(club_cdf, rng.random(crow.shape[0])), C - 1).astype(np.int64) + 1
    crow, cval = _dedup_keep_order(crow, cval)
    order = np.argsort(crow, kind="stable")
    crow, cval = crow[order], cval[order]
    clubs_ptr, clubs = _csr_from_rows(crow, cval.astype(np.uint32), n)

    # ---- text columns ---------------------------------------------------------------
    cell_cnt = np.zeros((n, T), np.int64)
    per_col = []
    fill_base = 0.10 + 0.55 * (comp / 100.0)        # per-user fill propensity tied to completion
    for t, name in enumerate(cols):
        crng = np.random.default_rng([seed & 0xFFFFFFFF, 7919 + t])
        if name in CATEGORICAL:
            V = int(crng.integers(8, 300))
            s = 1.05
            p_fill = np.clip(fill_base * 1.25, 0, 0.95)
            filled = crng.random(n) < p_fill
            k = np.where(crng.random(n) < 0.12, 2, 1)
        else:
            V = int(crng.integers(20_000, 300_000) * max(text_scale * min(1.0, n / 200_000), 0.02)) + 50
            s = 1.05
            p_fill = np.clip(fill_base * 0.85, 0, 0.95)
            filled = crng.random(n) < p_fill
            k = np.minimum(1 + crng.geometric(1 / 5.6, n), 16)
        k = np.where(filled, k, 0)
        rows = np.repeat(np.arange(n, dtype=np.int64), k)
        cdf = _zipf_sampler(V, s)
        toks = np.minimum(np.searchsorted(cdf, crng.random(rows.shape[0])), V - 1).astype(np.int64) + 1
        key = rows * (1 << 20) + toks
        ukey, cnt = np.unique(key, return_counts=True)        # sorted by (row, token)
        urow = ukey >> 20
        utok = (ukey & ((1 << 20) - 1)).astype(np.int32)
        tf = (cnt + (crng.random(cnt.shape[0]) < 0.1)).astype(np.int32)
        cell_cnt[:, t] = np.bincount(urow, minlength=n)
        per_col.append((urow, utok, tf))
    tok_ptr = np.zeros(n * T + 1, np.int64)
    np.cumsum(cell_cnt.reshape(-1), out=tok_ptr[1:])
    nnz = int(tok_ptr[-1])
    tok_id = np.zeros(nnz, np.int32)
    tok_tf = np.zeros(nnz, np.int32)
    for t, (urow, utok, tf) in enumerate(per_col):
        if urow.shape[0] == 0:
            continue
        start = tok_ptr[urow * T + t]
        # rank inside the cell: urow is sorted, so rank = i - first index of the run
        first = np.r_[0, np.flatnonzero(np.diff(urow)) + 1]
        run_len = np.diff(np.r_[first, urow.shape[0]])
        rank = np.arange(urow.shape[0]) - np.repeat(first, run_len)
        pos = start + rank
        tok_id[pos] = utok
        tok_tf[pos] = tf
    del per_col

    w = World(text_columns=cols, uid=uid, pub=pub, comp=comp, gender=gender, age=age, region=region,
              clubs_ptr=clubs_ptr, clubs=clubs, friends_ptr=friends_ptr, friends=friends,
              tok_ptr=tok_ptr, tok_id=tok_id, tok_tf=tok_tf, adj_uid=adj_uid, adj_ptr=adj_ptr, adj=adj)
    w.meta = dict(generator="synth.make_world", n=n, seed=int(seed), T=T, mean_deg=float(mean_deg),
                  edges=int(adj.shape[0]), mean_out_deg=float(adj.shape[0] / max(n, 1)),
                  tokens_per_user=float(nnz / max(n, 1)),
                  filled_cols_per_user=float((cell_cnt > 0).sum() / max(n, 1)),
                  clubs_per_user=float(clubs.shape[0] / max(n, 1)), n_clubs=int(C))
    return w


def sample_normalizers(world: World, seed=12345, pairs=4000):
    """Normalizers in the spirit of compute_column_normalizers (src/utils.cpp:155-240): mean and
    sample sd of the *raw* per-field similarities (zeros included, raw-count cosine for text) over
    random user pairs.  Not the reference's sampling (that one walks an unordered_map); these are
    *inputs* fed identically to every implementation, so only plausibility matters."""
    rng = np.random.default_rng(seed)
    n, T = world.n, world.T
    a = rng.integers(0, n, pairs)
    b = rng.integers(0, n, pairs)
    keep = a != b
    a, b = a[keep], b[keep]
    P = a.shape[0]
    vals = np.zeros((NUM_FIXED + T, P))
    w = world
    vals[0] = (w.pub[a] >= 0) & (w.pub[b] >= 0) & (w.pub[a] == w.pub[b])
    vals[1] = (w.gender[a] >= 0) & (w.gender[b] >= 0) & (w.gender[a] == w.gender[b])
    for row, arr in ((2, w.comp), (3, w.age)):
        x, y = arr[a].astype(np.float64), arr[b].astype(np.float64)
        ok = (x > 0) & (y > 0)
        vals[row] = np.where(ok, np.minimum(x, y) / np.maximum(np.maximum(x, y), 1), 0.0)
    ra, rb = w.region[a], w.region[b]
    ac, bc = (ra >= 0).sum(1), (rb >= 0).sum(1)
    m = ((ra >= 0) & (rb >= 0) & (ra == rb)).sum(1)
    vals[4] = np.where((ac > 0) & (bc > 0), m / np.sqrt(np.maximum(ac, 1) * np.maximum(bc, 1)), 0.0)
    for row, ptr, arr in ((5, w.clubs_ptr, w.clubs), (6, w.friends_ptr, w.friends)):
        for i in range(P):
            A = arr[ptr[a[i]]:ptr[a[i] + 1]]
            B = arr[ptr[b[i]]:ptr[b[i] + 1]]
            if A.size and B.size:
                vals[row, i] = np.isin(B, A).sum() / (np.sqrt(A.size) * np.sqrt(B.size))
    cell = np.diff(w.tok_ptr).reshape(n, T)
    both = (cell[a] > 0) & (cell[b] > 0)
    for i, t in zip(*np.nonzero(both)):
        ca, cb = a[i] * T + t, b[i] * T + t
        ia, ta = w.tok_id[w.tok_ptr[ca]:w.tok_ptr[ca + 1]], w.tok_tf[w.tok_ptr[ca]:w.tok_ptr[ca + 1]]
        ib, tb = w.tok_id[w.tok_ptr[cb]:w.tok_ptr[cb + 1]], w.tok_tf[w.tok_ptr[cb]:w.tok_ptr[cb + 1]]
        _, xa, xb = np.intersect1d(ia, ib, return_indices=True)
        if xa.size:
            dot = float((ta[xa].astype(np.float64) * tb[xb]).sum())
            vals[NUM_FIXED + t, i] = dot / (np.sqrt(float((ta.astype(np.float64) ** 2).sum()))
                                             * np.sqrt(float((tb.astype(np.float64) ** 2).sum())))
    mean = vals.mean(1)
    sd = vals.std(1, ddof=1) if P > 1 else np.ones(NUM_FIXED + T)
    sd[sd == 0] = 1.0
    return mean.astype(np.float32), sd.astype(np.float32)


def sample_queries(world: World, q, seed):
    rng = np.random.default_rng(seed)
    idx = rng.choice(world.n, size=min(q, world.n), replace=False)
    return world.uid[np.sort(idx)].astype(np.int32)


def normalizer_pair_sample(world: World, sample_size=100_000, comps_per_user=5, seed=12345):
    """The pair sample of the reference's compute_column_normalizers(profiles, cols, sample_size, comps_per_user)
    (src/utils.cpp:163-181): sample_size * comps_per_user DISTINCT unordered pairs (a != b) of users drawn uniformly
    with an MT19937 stream seeded 12345, at most 10 attempts per wanted pair.  The reference indexes its
    unordered_map's iteration order, which only its own C++ side can reproduce; here the users are indexed in
    world order, so the sample has the same distribution, not the same pairs.  Returns (a_uid, b_uid) in draw order
    -- the input of pk_normalizers_from_pairs."""
    n = world.n
    need = int(sample_size) * int(comps_per_user)
    need = min(need, n * (n - 1) // 2)
    rs = np.random.RandomState(seed)              # MT19937, like std::mt19937
    a_out, b_out = [], []
    seen = np.zeros(0, np.int64)
    have, attempts = 0, 0
    while have < need and attempts < need * 10:
        m = int((need - have) * 1.05) + 16
        a = rs.randint(0, n, m)
        b = rs.randint(0, n, m)
        attempts += m
        ok = a != b
        a, b = a[ok], b[ok]
        key = np.minimum(a, b).astype(np.int64) * n + np.maximum(a, b)
        _, first = np.unique(key, return_index=True)          # distinct inside the draw, first occurrence kept
        first.sort()
        a, b, key = a[first], b[first], key[first]
        if seen.size:
            new = ~np.isin(key, seen)
            a, b, key = a[new], b[new], key[new]
        take = min(len(a), need - have)
        a_out.append(a[:take]); b_out.append(b[:take])
        seen = np.concatenate([seen, key[:take]])
        have += take
    a = np.concatenate(a_out) if a_out else np.zeros(0, np.int64)
    b = np.concatenate(b_out) if b_out else np.zeros(0, np.int64)
    return world.uid[a].astype(np.int32), world.uid[b].astype(np.int32)
