"""Generates the golden fixtures of tests/golden/ by running the UNMODIFIED reference (compiled from
/root/reference into oracle/_ref/libpokec_ref.so by oracle/build_oracle.py) on a small seeded world.

    python tests/golden/make_golden.py

The reference cannot travel to the GPU box, the fixtures can: world_small.npz is the input (the flat form of
the reference's profiles / adj_list / normalizers), golden_small.npz holds what the reference's own public
API returned for it -- Recommender::profile_similarity, the per-field helpers, compute_idf_from_profiles,
the three recommend_* calls, and the full-scan loop over profile_similarity with the reference comparator.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from recommendation_system_pokec_b200.synth import make_world, sample_normalizers, sample_queries  # noqa: E402
from recommendation_system_pokec_b200.cabi import MODE_GRAPH, MODE_COLLAB, MODE_CLUBS  # noqa: E402
from oracle.oracle import CpuOracle, reference_available  # noqa: E402
from oracle import build_oracle  # noqa: E402


def golden_inputs(w):
    """the (deterministic) query / pair samples both the generator and the tests use"""
    rng = np.random.default_rng(2024)
    a = w.uid[rng.integers(0, w.n, 600)].astype(np.int32)
    b = w.uid[rng.integers(0, w.n, 600)].astype(np.int32)
    # half of the pairs: a user and one of its friends (non-trivial overlaps)
    rows = rng.integers(0, w.n, 300)
    for i, r in enumerate(rows):
        lo, hi = w.friends_ptr[r], w.friends_ptr[r + 1]
        if hi > lo:
            a[i] = w.uid[r]
            b[i] = np.int32(w.friends[lo + rng.integers(0, hi - lo)])
    q = sample_queries(w, 24, seed=77)
    return a, b, q


def main():
    build_oracle.build_all()
    assert reference_available(), "needs /root/reference (run in the build container)"
    w = make_world(n=400, seed=20240611)
    m, s = sample_normalizers(w, pairs=600)
    has = np.ones(w.n_fields, np.uint8)
    has[[4, 9, 30]] = 0                       # three fields without a normalizer entry: z = 6(s - 0.5)
    w.set_normalizers(m, s, has)
    w.save(os.path.join(HERE, "world_small.npz"))
    ref = CpuOracle(w, "reference")
    a, b, q = golden_inputs(w)
    out = {"pair_a": a, "pair_b": b, "queries": q,
           "pair_fas": ref.pair_similarity(a, b), "pair_fields": ref.pair_fields(a, b)}
    for col in (0, 2, 22, 32, 48):
        cells = np.arange(w.n) * w.T + col
        toks = np.unique(np.concatenate([w.tok_id[w.tok_ptr[c]:w.tok_ptr[c + 1]] for c in cells]))
        idf, present = ref.idf_lookup(col, toks)
        assert present.all() and ref.idf_size(col) == len(toks)
        out[f"idf_tok_{col}"] = toks.astype(np.int32)
        out[f"idf_val_{col}"] = idf
        idx, _ = ref.tfidf_index_lookup(col, toks)
        out[f"idf_dbl_{col}"] = idx
    for name, mode, limit in (("graph", MODE_GRAPH, 5000), ("graph_l25", MODE_GRAPH, 25), ("collab", MODE_COLLAB, 5000),
                              ("collab_l40", MODE_COLLAB, 40), ("clubs", MODE_CLUBS, 5000)):
        ids, sc, cnt = ref.recommend(mode, q, 20, limit)
        out[f"{name}_id"], out[f"{name}_score"], out[f"{name}_cnt"] = ids, sc, cnt
    for ex in (0, 1, 2):
        ids, sc, cnt = ref.fullscan(q, 30, ex)
        out[f"fullscan{ex}_id"], out[f"fullscan{ex}_score"], out[f"fullscan{ex}_cnt"] = ids, sc, cnt
    np.savez_compressed(os.path.join(HERE, "golden_small.npz"), **out)
    print("wrote", os.path.join(HERE, "golden_small.npz"), {k: v.shape for k, v in out.items() if k.endswith("_id")})


if __name__ == "__main__":
    main()
