"""Diagnostic (GPU box): pre-filter error statistics and two-stage vs exact full-scan agreement."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from recommendation_system_pokec_b200.synth import make_world, sample_normalizers, sample_queries
from recommendation_system_pokec_b200.engine import DeviceRecommender
from recommendation_system_pokec_b200 import engine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
w = make_world(n=n, seed=7)
m, s = sample_normalizers(w, pairs=1500)
w.set_normalizers(m, s)
dev = DeviceRecommender(w)
q = sample_queries(w, 32, seed=3)
a, e = dev.debug_prefilter(q, 0, min(w.n, 20000))
d = np.abs(a.astype(np.float64) - e)
print("margin", dev.stat(engine.STAT_MARGIN_E9) / 1e9, "max err", d.max(), "mean err", d.mean())
r, c = np.unravel_index(np.argmax(d), d.shape)
print("worst pair: row", r, "uid", w.uid[r], "query", q[c], "approx", a[r, c], "exact", e[r, c])
big = np.argwhere(d > 1e-4)
print("pairs with err > 1e-4:", len(big), "of", d.size)
for r, c in big[:10]:
    print("   row", r, "q", q[c], "approx", a[r, c], "exact", e[r, c])
for excl in (0, 1, 2):
    os.environ.pop("POKEC_B200_NO_PREFILTER", None)
    q2 = sample_queries(w, 70, seed=3)
    two = dev.fullscan(q2, 50, excl)
    fb = dev.stat(engine.STAT_FALLBACK_QUERIES)
    os.environ["POKEC_B200_NO_PREFILTER"] = "1"
    ex = dev.fullscan(q2, 50, excl)
    bad = np.argwhere(two[0] != ex[0])
    print(f"exclude={excl}: id mismatches {len(bad)}, fallback queries {fb}, first {bad[:5].tolist()}")
    for qq, j in bad[:3]:
        print("   q", q2[qq], "rank", j, "two-stage", two[0][qq, j], two[1][qq, j], "exact", ex[0][qq, j], ex[1][qq, j])
