"""The reference's friends hold-out harness (src/test.cpp:13-105, called from src/main.cpp:147) run UNMODIFIED on both
sides: the compiled reference and the CUDA-backed drop-in `Recommender`.  The harness builds its own Recommender over a
copy of the adjacency, copies the base recommender's public normalizer / idf members into it, removes a fifth of each
tested user's friends BEHIND the live object (the rows stay shortened for the users tested later) and asks
recommend_collaborative(uid, F/5, 1000).  Its output file -- one hit ratio per tested user -- must be identical.

src/test.cpp is compiled where it lies into both test libraries (oracle/build_oracle.py, dropin/build_dropin.py)."""
import os
import time

import pytest

from tests.conftest import make_test_world

pytestmark = pytest.mark.gpu


def test_reference_holdout_harness_identical_on_the_dropin(oracle_libs, tmp_path):
    from oracle.oracle import CpuOracle, reference_available, dropin_available
    if not (dropin_available() and reference_available()):
        pytest.skip("needs oracle/_ref/libpokec_ref.so and dropin/_build/libpokec_dropin.so (built where the reference is)")
    w = make_test_world(n=3000, seed=77, mean_deg=30.0)
    ref = CpuOracle(w, "reference")
    drop = CpuOracle(w, "dropin")
    if not hasattr(ref.lib, "ref_holdout_test") or not hasattr(drop.lib, "ref_holdout_test"):
        pytest.skip("test libraries predate ref_holdout_test")
    t0 = time.perf_counter()
    r = ref.holdout_test(60, os.path.join(tmp_path, "ref.csv"))
    t1 = time.perf_counter()
    d = drop.holdout_test(60, os.path.join(tmp_path, "drop.csv"))
    t2 = time.perf_counter()
    ref.close()
    drop.close()
    assert len(r) == 60, "the world must hold enough users with >= 20 friends"
    assert d == r
    assert any(float(x) > 0 for x in r), "a harness that never hits would compare nothing"
    print(f"hold-out harness: {len(r)} users, reference {1e3 * (t1 - t0) / len(r):.2f} ms/user, "
          f"drop-in {1e3 * (t2 - t1) / len(r):.2f} ms/user")
