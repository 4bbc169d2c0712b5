"""ctypes binding of libpokec_b200.so (include/pokec_b200.h) -- the Python-side handle of the device store.

`DeviceRecommender` mirrors the reference's C++ `Recommender` (include/recommender.h:17-71) method by
method, batched: recommend_graph_registration / recommend_by_interest / recommend_collaborative /
recommend_clubs_collab, set_*_normalizers, compute_idf_from_profiles, set_tfidf_index,
profile_similarity -- plus the north-star full scan.  It fails loudly when the CUDA library or a
device is missing; there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

from .cabi import (PkWorld, make_pk_world, ptr, i32p, f32p, u8p, MODE_GRAPH, MODE_COLLAB, MODE_CLUBS,
                   EXCLUDE_NONE, EXCLUDE_SELF, EXCLUDE_SELF_AND_ADJ)
from .world import World

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libpokec_b200.so")

# every symbol include/pokec_b200.h declares
ABI_SYMBOLS = (
    "pk_last_error", "pk_abi_version", "pk_device_count", "pk_store_create", "pk_store_destroy",
    "pk_store_load_world", "pk_store_set_shard", "pk_set_normalizers", "pk_compute_idf", "pk_set_idf",
    "pk_idf_size", "pk_get_idf", "pk_set_adj_row", "pk_pair_similarity", "pk_pair_fields", "pk_pair_overlap",
    "pk_fullscan_topk", "pk_fullscan_prepare", "pk_fullscan_execute", "pk_fullscan_fetch", "pk_fullscan_keys", "pk_recommend", "pk_candidates", "pk_fullscan_local_keys", "pk_merge_keys", "pk_sync",
    "pk_normalizers_from_pairs", "pk_edge_cache_build", "pk_edge_cache_drop", "pk_edge_cache_row", "pk_fullscan_execute_shortlists", "pk_fullscan_shortlist_keys", "pk_fullscan_execute_finish",
    "pk_stream", "pk_stat", "pk_debug_prefilter", "pk_world_save", "pk_world_open", "pk_world_close", "pk_candidates_batch", "pk_candidates_batch_noalloc", "pk_store_set_option", "pk_fullscan_theta_export", "pk_fullscan_theta_ptr", "pk_fullscan_theta_peers",
)
STAT_KERNEL_LAUNCHES, STAT_PAIRS_SCORED, STAT_STORE_BYTES, STAT_SCAN_KERNEL_NS, STAT_SCAN_LAUNCHES, STAT_GROUPS, \
    STAT_SHARD_RECORD_BYTES, STAT_H2D_BYTES, STAT_D2H_BYTES, STAT_PREFILTER, STAT_FALLBACK_QUERIES, STAT_RESCORE_NS, \
    STAT_SHARD_LEAN_BYTES, STAT_SHORTLIST, STAT_MARGIN_E9, STAT_EDGE_CACHE, STAT_PREP_NS, STAT_MERGE_NS = range(18)

OPT_PREFILTER_RESERVE_SMS = 1

_lib = None


class PokecError(RuntimeError):
    pass


def load_library():
    """dlopen the in-tree CUDA library.  No fallback: a missing build is an error."""
    global _lib
    if _lib is None:
        # POKEC_B200_LIB: another build of the same CUDA library (kernel A/B experiments, tools/variants.py)
        path = os.environ.get("POKEC_B200_LIB") or LIB_PATH
        if not os.path.exists(path):
            raise PokecError(f"{path} is missing: run `python -m recommendation_system_pokec_b200.build` "
                             "(there is no CPU implementation of the scoring path)")
        lib = C.CDLL(path)
        lib.pk_last_error.restype = C.c_char_p
        lib.pk_stat.restype = C.c_int64
        lib.pk_stat.argtypes = [C.c_void_p, C.c_int]
        lib.pk_stream.restype = C.c_void_p
        lib.pk_stream.argtypes = [C.c_void_p]
        lib.pk_store_destroy.argtypes = [C.c_void_p]
        lib.pk_store_destroy.restype = None
        lib.pk_world_close.argtypes = [C.c_void_p]
        lib.pk_world_close.restype = None
        _lib = lib
    return _lib


def _check(lib, rc):
    if rc != 0:
        raise PokecError(f"libpokec_b200 error {rc}: {lib.pk_last_error().decode()}")


def world_save(world: World, path: str):
    """the world as one mmap-able binary file (pk_world_save; no GPU needed)"""
    lib = load_library()
    pw = make_pk_world(world)
    _check(lib, lib.pk_world_save(C.byref(pw), os.fsencode(path)))


class WorldFile:
    """A world file mapped read-only (pk_world_open).  `view` is the pk_world whose pointers point into the mapping;
    `to_world()` copies it into a `World` (column names are not part of the file: pass them or get c0, c1, ...)."""

    def __init__(self, path: str):
        self.lib = load_library()
        self.h = C.c_void_p()
        self.view = PkWorld()
        _check(self.lib, self.lib.pk_world_open(os.fsencode(path), C.byref(self.h), C.byref(self.view)))

    def close(self):
        if getattr(self, "h", None):
            self.lib.pk_world_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def to_world(self, text_columns=None) -> World:
        v = self.view
        n, T, m = int(v.n_users), int(v.n_text_cols), int(v.n_adj_rows)

        def arr(p, count, dtype):
            if count == 0 or not p:
                return np.zeros(count, dtype)
            return np.ctypeslib.as_array(p, shape=(count,)).astype(dtype, copy=True)

        cp, fp, tp = arr(v.clubs_ptr, n + 1, np.int64), arr(v.friends_ptr, n + 1, np.int64), arr(v.tok_ptr, n * T + 1, np.int64)
        ap = arr(v.adj_ptr, m + 1, np.int64)
        ntk = int(tp[-1])
        return World(text_columns=list(text_columns) if text_columns is not None else [f"c{i}" for i in range(T)],
                     uid=arr(v.uid, n, np.int32), pub=arr(v.pub, n, np.int32), comp=arr(v.comp, n, np.int32),
                     gender=arr(v.gender, n, np.int32), age=arr(v.age, n, np.int32),
                     region=arr(v.region, 3 * n, np.int32).reshape(n, 3),
                     clubs_ptr=cp, clubs=arr(v.clubs, int(cp[-1]), np.uint32),
                     friends_ptr=fp, friends=arr(v.friends, int(fp[-1]), np.uint32),
                     tok_ptr=tp, tok_id=arr(v.tok_id, ntk, np.int32), tok_tf=arr(v.tok_tf, ntk, np.int32),
                     adj_uid=arr(v.adj_uid, m, np.int32), adj_ptr=ap, adj=arr(v.adj, int(ap[-1]), np.int32),
                     norm_mean=arr(v.norm_mean, 7 + T, np.float32), norm_sd=arr(v.norm_sd, 7 + T, np.float32),
                     norm_has=arr(v.norm_has, 7 + T, np.uint8))


class DeviceRecommender:
    def __init__(self, world: World, device: int = 0, compute_idf: bool = True, _view=None):
        self.lib = load_library()
        self.world = world
        self.h = C.c_void_p()
        _check(self.lib, self.lib.pk_store_create(int(device), C.byref(self.h)))
        self._pw = _view if _view is not None else make_pk_world(world)
        _check(self.lib, self.lib.pk_store_load_world(self.h, C.byref(self._pw)))
        if compute_idf:
            self.compute_idf()

    @classmethod
    def from_world_file(cls, path: str, device: int = 0, compute_idf: bool = True, text_columns=None):
        """builds the device store straight from the mapped file (pk_world_open -> pk_store_load_world: no parse and
        no host copy of the big arrays); `world` holds only what the host-side helpers need (sizes, names)"""
        import types
        with WorldFile(path) as f:
            T, n = int(f.view.n_text_cols), int(f.view.n_users)
            names = list(text_columns) if text_columns is not None else [f"c{i}" for i in range(T)]
            shape = types.SimpleNamespace(T=T, n=n, n_fields=7 + T, text_columns=names)   # what the methods read from self.world
            return cls(shape, device=device, compute_idf=compute_idf, _view=f.view)

    def close(self):
        if getattr(self, "h", None):
            self.lib.pk_store_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- setters (reference src/recommender.cpp:27-66) ---------------------------------------
    def set_normalizers(self, mean, sd, has=None):
        mean = np.ascontiguousarray(mean, np.float32)
        sd = np.ascontiguousarray(sd, np.float32)
        has_p = None if has is None else ptr(np.ascontiguousarray(has, np.uint8), u8p)
        _check(self.lib, self.lib.pk_set_normalizers(self.h, ptr(mean, f32p), ptr(sd, f32p), has_p, len(mean)))

    def compute_idf(self, double_log=False):
        _check(self.lib, self.lib.pk_compute_idf(self.h, int(double_log)))

    def set_idf(self, col, tokens, idf):
        t = np.ascontiguousarray(tokens, np.int32)
        v = np.ascontiguousarray(idf, np.float32)
        assert t.shape == v.shape
        _check(self.lib, self.lib.pk_set_idf(self.h, int(col), ptr(t, i32p), ptr(v, f32p), C.c_int64(len(t))))

    def erase_idf_column(self, col):
        _check(self.lib, self.lib.pk_set_idf(self.h, int(col), None, None, C.c_int64(-1)))

    def idf_size(self, col):
        n = C.c_int64()
        _check(self.lib, self.lib.pk_idf_size(self.h, int(col), C.byref(n)))
        return int(n.value)

    def get_idf(self, col):
        n = self.idf_size(col)
        if n < 0:
            return None
        tok = np.zeros(max(n, 1), np.int32)
        idf = np.zeros(max(n, 1), np.float32)
        df = np.zeros(max(n, 1), np.int32)
        _check(self.lib, self.lib.pk_get_idf(self.h, int(col), ptr(tok, i32p), ptr(idf, f32p), ptr(df, i32p), C.c_int64(n)))
        return tok[:n], idf[:n], df[:n]

    def erase_idf_tokens(self, col, tokens):
        tok, idf, _ = self.get_idf(col)
        keep = ~np.isin(tok, np.asarray(tokens, np.int32))
        self.set_idf(col, tok[keep], idf[keep])

    def idf_lookup(self, col, tokens):
        got = self.get_idf(col)
        t = np.asarray(tokens, np.int32)
        if got is None:
            return np.zeros(len(t), np.float32), np.zeros(len(t), bool)
        tok, idf, _ = got
        pos = np.searchsorted(tok, t)
        pos_c = np.minimum(pos, max(len(tok) - 1, 0))
        has = (pos < len(tok)) & (tok[pos_c] == t) if len(tok) else np.zeros(len(t), bool)
        return np.where(has, idf[pos_c] if len(tok) else 0, 0).astype(np.float32), has

    def set_adj_row(self, uid, nb):
        nb = np.ascontiguousarray(nb, np.int32)
        _check(self.lib, self.lib.pk_set_adj_row(self.h, int(uid), ptr(nb, i32p), C.c_int64(len(nb))))

    def set_option(self, which, value):
        _check(self.lib, self.lib.pk_store_set_option(self.h, int(which), C.c_int64(int(value))))

    def set_shard(self, row_begin, row_end):
        _check(self.lib, self.lib.pk_store_set_shard(self.h, C.c_int64(row_begin), C.c_int64(row_end)))

    # ---- pairs -----------------------------------------------------------------------------------
    def pair_similarity(self, a, b):
        a = np.ascontiguousarray(a, np.int32)
        b = np.ascontiguousarray(b, np.int32)
        out = np.zeros(len(a), np.float32)
        _check(self.lib, self.lib.pk_pair_similarity(self.h, C.c_int64(len(a)), ptr(a, i32p), ptr(b, i32p), ptr(out, f32p)))
        return out

    def pair_fields(self, a, b):
        a = np.ascontiguousarray(a, np.int32)
        b = np.ascontiguousarray(b, np.int32)
        out = np.zeros((len(a), 3 + self.world.T), np.float32)
        _check(self.lib, self.lib.pk_pair_fields(self.h, C.c_int64(len(a)), ptr(a, i32p), ptr(b, i32p), ptr(out, f32p)))
        return out

    def normalizers_from_pairs(self, a, b):
        """compute_column_normalizers (src/utils.cpp:155-240) on a given pair sample -> (mean, sd), 7 + T floats each"""
        a = np.ascontiguousarray(a, np.int32)
        b = np.ascontiguousarray(b, np.int32)
        nf = self.world.n_fields
        mean, sd = np.zeros(nf, np.float32), np.zeros(nf, np.float32)
        _check(self.lib, self.lib.pk_normalizers_from_pairs(self.h, C.c_int64(len(a)), ptr(a, i32p), ptr(b, i32p),
                                                            ptr(mean, f32p), ptr(sd, f32p)))
        return mean, sd

    # ---- per-edge FAS cache (clubs) ---------------------------------------------------------------
    def edge_cache_build(self):
        _check(self.lib, self.lib.pk_edge_cache_build(self.h))

    def edge_cache_drop(self):
        _check(self.lib, self.lib.pk_edge_cache_drop(self.h))

    def edge_cache_row(self, uid, cap=1 << 16):
        out = np.zeros(cap, np.float32)
        n = C.c_int64()
        _check(self.lib, self.lib.pk_edge_cache_row(self.h, int(uid), ptr(out, f32p), C.c_int64(cap), C.byref(n)))
        return out[:min(int(n.value), cap)].copy()

    def pair_overlap(self, a, b, which):
        a = np.ascontiguousarray(a, np.int32)
        b = np.ascontiguousarray(b, np.int32)
        out = np.zeros(len(a), np.int32)
        _check(self.lib, self.lib.pk_pair_overlap(self.h, C.c_int64(len(a)), ptr(a, i32p), ptr(b, i32p), int(which), ptr(out, i32p)))
        return out

    # ---- recommendations ---------------------------------------------------------------------------
    @staticmethod
    def _alloc(nq, k):
        return (np.full((nq, max(k, 0)), -1, np.int32), np.zeros((nq, max(k, 0)), np.float32), np.zeros(nq, np.int32))

    def fullscan(self, q, k, exclude_mode=EXCLUDE_SELF_AND_ADJ):
        q = np.ascontiguousarray(q, np.int32)
        ids, sc, cnt = self._alloc(len(q), k)
        _check(self.lib, self.lib.pk_fullscan_topk(self.h, len(q), ptr(q, i32p), int(k), int(exclude_mode),
                                                   ptr(ids, i32p), ptr(sc, f32p), ptr(cnt, i32p)))
        return ids, sc, cnt

    def fullscan_prepare(self, q, k, exclude_mode=EXCLUDE_SELF_AND_ADJ):
        q = np.ascontiguousarray(q, np.int32)
        self._job_shape = (len(q), k)
        _check(self.lib, self.lib.pk_fullscan_prepare(self.h, len(q), ptr(q, i32p), int(k), int(exclude_mode)))

    def fullscan_execute(self):
        _check(self.lib, self.lib.pk_fullscan_execute(self.h))

    def theta_export(self) -> bytes:
        """the cudaIpc handle of this store's pruning-threshold array (64 bytes)"""
        buf = C.create_string_buffer(64)
        _check(self.lib, self.lib.pk_fullscan_theta_export(self.h, buf))
        return buf.raw

    def theta_ptr(self) -> int:
        self.lib.pk_fullscan_theta_ptr.restype = C.c_void_p
        self.lib.pk_fullscan_theta_ptr.argtypes = [C.c_void_p]
        return int(self.lib.pk_fullscan_theta_ptr(self.h) or 0)

    def theta_peers(self, handles: bytes = None, n_ranks: int = 1, rank: int = 0, ptrs=None):
        """maps the other ranks' threshold arrays (handles: n_ranks x 64 bytes) or, in one process, their device pointers"""
        if ptrs is not None:
            arr = (C.c_void_p * n_ranks)(*[C.c_void_p(p) for p in ptrs])
            _check(self.lib, self.lib.pk_fullscan_theta_peers(self.h, None, int(n_ranks), int(rank), arr))
        else:
            _check(self.lib, self.lib.pk_fullscan_theta_peers(self.h, handles, int(n_ranks), int(rank), None))

    def fullscan_execute_shortlists(self):
        _check(self.lib, self.lib.pk_fullscan_execute_shortlists(self.h))

    def fullscan_shortlist_keys(self, d_short_ptr: int):
        _check(self.lib, self.lib.pk_fullscan_shortlist_keys(self.h, C.c_void_p(d_short_ptr)))

    def fullscan_execute_finish(self, d_all_short_ptr, n_lists, rank, n_ranks):
        p = C.c_void_p(d_all_short_ptr) if d_all_short_ptr else None
        _check(self.lib, self.lib.pk_fullscan_execute_finish(self.h, p, int(n_lists), int(rank), int(n_ranks)))

    def fullscan_fetch(self):
        ids, sc, cnt = self._alloc(*self._job_shape)
        _check(self.lib, self.lib.pk_fullscan_fetch(self.h, ptr(ids, i32p), ptr(sc, f32p), ptr(cnt, i32p)))
        return ids, sc, cnt

    def fullscan_keys(self, d_keys_ptr: int):
        _check(self.lib, self.lib.pk_fullscan_keys(self.h, C.c_void_p(d_keys_ptr)))

    def recommend(self, mode, q, k, limit=10000):
        q = np.ascontiguousarray(q, np.int32)
        ids, sc, cnt = self._alloc(len(q), k)
        _check(self.lib, self.lib.pk_recommend(self.h, int(mode), len(q), ptr(q, i32p), int(k), int(limit),
                                               ptr(ids, i32p), ptr(sc, f32p), ptr(cnt, i32p)))
        return ids, sc, cnt

    # reference method names (include/recommender.h:24-35), batched over users
    def recommend_graph_registration(self, users, topk, candidate_limit=10000):
        return self.recommend(MODE_GRAPH, users, topk, candidate_limit)

    recommend_by_interest = recommend_graph_registration
    recommend_friends_graph = recommend_graph_registration
    recommend_friends_by_interest = recommend_graph_registration

    def recommend_collaborative(self, users, topk, candidate_limit=10000):
        return self.recommend(MODE_COLLAB, users, topk, candidate_limit)

    recommend_friends_collab = recommend_collaborative

    def recommend_clubs_collab(self, users, topk, candidate_limit=10000):
        return self.recommend(MODE_CLUBS, users, topk, candidate_limit)

    def candidates(self, mode, user, limit, cap=1 << 20):
        out = np.zeros(cap, np.int32)
        n = C.c_int64()
        _check(self.lib, self.lib.pk_candidates(self.h, int(mode), int(user), int(limit), ptr(out, i32p), C.c_int64(cap), C.byref(n)))
        return out[:min(int(n.value), cap)].copy()

    def candidates_batch(self, mode, users, limit):
        """raw candidate lists of many users, expanded on the device -> (ids [nq][limit], counts [nq])"""
        q = np.ascontiguousarray(users, np.int32)
        ids = np.full((len(q), int(limit)), -1, np.int32)
        cnt = np.zeros(len(q), np.int32)
        _check(self.lib, self.lib.pk_candidates_batch(self.h, int(mode), len(q), ptr(q, i32p), int(limit), ptr(ids, i32p), ptr(cnt, i32p)))
        return ids, cnt

    def candidates_batch_noalloc(self, mode, d_q_uid_ptr: int, nq, limit, d_out_ids_ptr: int, d_out_cnt_ptr: int):
        """pk_candidates_batch on DEVICE buffers with static scratch (no allocation inside the call; limit <= 992,
        nq <= 8192; the adjacency must already be resident).  Asynchronous on the store's stream: sync() before reading."""
        _check(self.lib, self.lib.pk_candidates_batch_noalloc(self.h, int(mode), int(nq), C.c_void_p(d_q_uid_ptr), int(limit),
                                                              C.c_void_p(d_out_ids_ptr), C.c_void_p(d_out_cnt_ptr)))

    # ---- multi-GPU plumbing --------------------------------------------------------------------------
    def fullscan_local_keys(self, q, k, exclude_mode, d_keys_ptr: int):
        q = np.ascontiguousarray(q, np.int32)
        _check(self.lib, self.lib.pk_fullscan_local_keys(self.h, len(q), ptr(q, i32p), int(k), int(exclude_mode),
                                                         C.c_void_p(d_keys_ptr)))

    def merge_keys(self, d_keys_ptr: int, n_lists, nq, k, fetch=True):
        if not fetch:      # device merge only, asynchronous
            _check(self.lib, self.lib.pk_merge_keys(self.h, C.c_void_p(d_keys_ptr), int(n_lists), int(nq), int(k),
                                                    None, None, None))
            return None
        ids, sc, cnt = self._alloc(nq, k)
        _check(self.lib, self.lib.pk_merge_keys(self.h, C.c_void_p(d_keys_ptr), int(n_lists), int(nq), int(k),
                                                ptr(ids, i32p), ptr(sc, f32p), ptr(cnt, i32p)))
        return ids, sc, cnt

    def debug_prefilter(self, q, row_begin, row_end):
        """(approx, exact) FAS of <= 32 queries against store rows [row_begin,row_end), each [rows][nq]"""
        q = np.ascontiguousarray(q, np.int32)
        rows = int(row_end - row_begin)
        a = np.zeros((rows, len(q)), np.float32)
        e = np.zeros((rows, len(q)), np.float32)
        _check(self.lib, self.lib.pk_debug_prefilter(self.h, len(q), ptr(q, i32p), C.c_int64(row_begin), C.c_int64(row_end),
                                                     ptr(a, f32p), ptr(e, f32p)))
        return a, e

    def sync(self):
        _check(self.lib, self.lib.pk_sync(self.h))

    def stream(self) -> int:
        return int(self.lib.pk_stream(self.h) or 0)

    def stat(self, which) -> int:
        return int(self.lib.pk_stat(self.h, int(which)))
