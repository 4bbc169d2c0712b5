"""ctypes front-ends of the two checkers (TEST INFRASTRUCTURE, see oracle/fas_oracle.cpp header):

  CpuOracle(world, kind="port")       oracle/libpokec_oracle.so   -- the CPU restatement
  CpuOracle(world, kind="reference")  oracle/_ref/libpokec_ref.so -- the compiled, unmodified reference

Both expose the same methods.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

from recommendation_system_pokec_b200.cabi import (PkWorld, make_pk_world, col_names_array, ptr, i32p, f32p, u8p)
from recommendation_system_pokec_b200.world import World

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libpokec_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libpokec_ref.so")
# the CUDA-backed drop-in `Recommender` behind the same ref_* driver (dropin/build_dropin.py); it is the
# PRODUCT under test, loaded through this front-end only because it shares the driver's entry points
DROPIN_SO = os.path.join(os.path.dirname(HERE), "dropin", "_build", "libpokec_dropin.so")


def reference_available() -> bool:
    return os.path.exists(REF_SO)


def dropin_available() -> bool:
    return os.path.exists(DROPIN_SO)


def ensure_built():
    from . import build_oracle
    build_oracle.build_all()


class CpuOracle:
    def __init__(self, world: World, kind: str = "port", compute_idf: bool = True):
        assert kind in ("port", "reference", "dropin")
        self.dropin = kind == "dropin"
        if self.dropin:
            kind = "reference"          # same driver, same entry points
        self.kind = kind
        self.world = world
        self.p = "orc_" if kind == "port" else "ref_"
        path = DROPIN_SO if self.dropin else (PORT_SO if kind == "port" else REF_SO)
        if not os.path.exists(path) and not self.dropin:
            ensure_built()
        self.lib = C.CDLL(path)
        self._pw = make_pk_world(world)
        create = self._fn("create")
        create.restype = C.c_void_p
        if kind == "port":
            self.h = C.c_void_p(create(C.byref(self._pw)))
        else:
            self._names = col_names_array(world)
            self.h = C.c_void_p(create(C.byref(self._pw), self._names))
        if compute_idf:
            self.compute_idf()

    def _fn(self, name):
        return getattr(self.lib, self.p + name)

    def close(self):
        if self.h:
            self._fn("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- idf -------------------------------------------------------------------------------
    def compute_idf(self, double_log=False):
        if self.kind == "port":
            self._fn("compute_idf")(self.h, int(double_log))
        else:
            assert not double_log
            self._fn("compute_idf")(self.h)

    def erase_idf_column(self, col):
        self._fn("erase_idf_column")(self.h, int(col))

    def erase_idf_tokens(self, col, tokens):
        t = np.ascontiguousarray(tokens, np.int32)
        self._fn("erase_idf_tokens")(self.h, int(col), len(t), ptr(t, i32p))

    def idf_lookup(self, col, tokens):
        t = np.ascontiguousarray(tokens, np.int32)
        out = np.zeros(len(t), np.float32)
        has = np.zeros(len(t), np.uint8)
        if self.kind == "port":
            self._fn("idf_lookup")(self.h, int(col), len(t), ptr(t, i32p), ptr(out, f32p), ptr(has, u8p), None)
        else:
            self._fn("idf_lookup")(self.h, int(col), len(t), ptr(t, i32p), ptr(out, f32p), ptr(has, u8p))
        return out, has.astype(bool)

    def idf_size(self, col):
        f = self._fn("idf_size")
        f.restype = C.c_int64
        return int(f(self.h, int(col)))

    def tfidf_index_lookup(self, col, tokens):
        """TFIDFIndex::build variant (double log) -- reference kind only."""
        assert self.kind == "reference"
        t = np.ascontiguousarray(tokens, np.int32)
        out = np.zeros(len(t), np.float32)
        has = np.zeros(len(t), np.uint8)
        self.lib.ref_tfidf_index_lookup(self.h, int(col), len(t), ptr(t, i32p), ptr(out, f32p), ptr(has, u8p))
        return out, has.astype(bool)

    # ---- pairs -----------------------------------------------------------------------------
    def pair_similarity(self, a, b, nthreads=1):
        a = np.ascontiguousarray(a, np.int32)
        b = np.ascontiguousarray(b, np.int32)
        out = np.zeros(len(a), np.float32)
        self._fn("pair_similarity")(self.h, C.c_int64(len(a)), ptr(a, i32p), ptr(b, i32p), ptr(out, f32p),
                                    int(nthreads))
        return out

    def pair_fields(self, a, b, nthreads=1):
        a = np.ascontiguousarray(a, np.int32)
        b = np.ascontiguousarray(b, np.int32)
        out = np.zeros((len(a), 3 + self.world.T), np.float32)
        self._fn("pair_fields")(self.h, C.c_int64(len(a)), ptr(a, i32p), ptr(b, i32p), ptr(out, f32p),
                                int(nthreads))
        return out

    def pair_overlap(self, a, b, which):
        assert self.kind == "port"
        a = np.ascontiguousarray(a, np.int32)
        b = np.ascontiguousarray(b, np.int32)
        out = np.zeros(len(a), np.int32)
        self.lib.orc_pair_overlap(self.h, C.c_int64(len(a)), ptr(a, i32p), ptr(b, i32p), int(which), ptr(out, i32p))
        return out

    # ---- recommendations ---------------------------------------------------------------------
    def _alloc(self, nq, k):
        return (np.full((nq, max(k, 0)), -1, np.int32), np.zeros((nq, max(k, 0)), np.float32),
                np.zeros(nq, np.int32))

    def recommend(self, mode, q, k, limit=10000, nthreads=1):
        q = np.ascontiguousarray(q, np.int32)
        ids, sc, cnt = self._alloc(len(q), k)
        self._fn("recommend")(self.h, int(mode), len(q), ptr(q, i32p), int(k), int(limit), ptr(ids, i32p),
                              ptr(sc, f32p), ptr(cnt, i32p), int(nthreads))
        return ids, sc, cnt

    def fullscan(self, q, k, exclude_mode=2, nthreads=1):
        q = np.ascontiguousarray(q, np.int32)
        ids, sc, cnt = self._alloc(len(q), k)
        self._fn("fullscan")(self.h, len(q), ptr(q, i32p), int(k), int(exclude_mode), ptr(ids, i32p),
                             ptr(sc, f32p), ptr(cnt, i32p), int(nthreads))
        return ids, sc, cnt

    def candidates(self, mode, user, limit, cap=1 << 20):
        assert self.kind == "port"
        out = np.zeros(cap, np.int32)
        n = self.lib.orc_candidates(self.h, int(mode), int(user), int(limit), ptr(out, i32p), int(cap))
        return out[:min(n, cap)].copy()

    def scores_of_query(self, quid, nthreads=1):
        assert self.kind == "port"
        out = np.zeros(self.world.n, np.float32)
        rc = self.lib.orc_scores_of_query(self.h, int(quid), ptr(out, f32p), int(nthreads))
        assert rc == 0
        return out

    def column_normalizers(self, sample_size=100000, comps_per_user=5):
        """the reference's compute_column_normalizers (src/utils.cpp:155-240) -- reference kind only."""
        assert self.kind == "reference"
        nf = self.world.n_fields
        mean = np.zeros(nf, np.float32)
        sd = np.zeros(nf, np.float32)
        self.lib.ref_column_normalizers(self.h, int(sample_size), int(comps_per_user), ptr(mean, f32p), ptr(sd, f32p))
        return mean, sd

    def normalizer_pairs(self, sample_size=100000, comps_per_user=5):
        """the (a, b) uid pairs compute_column_normalizers samples (src/utils.cpp:163-181), in its order"""
        assert self.kind == "reference"
        cap = int(sample_size) * int(comps_per_user)
        a = np.zeros(max(cap, 1), np.int32)
        b = np.zeros(max(cap, 1), np.int32)
        self.lib.ref_normalizer_pairs.restype = C.c_longlong
        n = int(self.lib.ref_normalizer_pairs(self.h, int(sample_size), int(comps_per_user), ptr(a, i32p), ptr(b, i32p),
                                              C.c_longlong(cap)))
        return a[:n].copy(), b[:n].copy()

    def holdout_test(self, sample_size, out_path):
        """the reference's friends hold-out harness (src/test.cpp:13-105), unmodified, on this handle's Recommender (the
        compiled reference, or the CUDA-backed drop-in); returns the ratios it wrote, one per tested user"""
        assert self.kind == "reference"
        self.lib.ref_holdout_test(self.h, int(sample_size), str(out_path).encode())
        with open(out_path) as f:
            return [line.strip() for line in f if line.strip()]

    def set_adj_row(self, uid, nb):
        assert self.kind == "reference"
        nb = np.ascontiguousarray(nb, np.int32)
        self.lib.ref_set_adj_row(self.h, int(uid), len(nb), ptr(nb, i32p))
