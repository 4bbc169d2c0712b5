"""Flat ("world") representation of the reference's in-memory model.

The reference keeps `unordered_map<int, UserProfile>` + `unordered_map<int, vector<int>>`
(include/user_profile.h:10-20, include/recommender.h:46-49).  Everything on our side --
the CUDA store, the CPU oracle, the compiled-reference driver and the synthetic
generator -- exchanges the same data as plain arrays:

  uid[n]                      int32   profile key (any int, unique)
  pub/comp/gender/age[n]      int32   UserProfile scalar fields (-1 / 0 = missing as in the reference)
  region[n,3]                 int32   region_parts, -1 = missing
  clubs_ptr[n+1], clubs[]     int64 / uint32   UserProfile::clubs   (order preserved)
  friends_ptr[n+1], friends[] int64 / uint32   UserProfile::friends (order preserved)
  tok_ptr[n*T+1], tok_id[], tok_tf[]   int64 / int32 / int32
                              token_cols as cells, row-major (user, column); ids unique per cell
  adj_uid[m], adj_ptr[m+1], adj[]      int32 / int64 / int32
                              adj_list rows in original neighbour order; rows may exist for
                              uids without a profile and neighbours may lack profiles
  norm_mean/sd[7+T] float32, norm_has[7+T] uint8
                              field_normalizers (7 fixed keys, order public, gender, completion,
                              age, region, clubs, friends) followed by column_normalizers
"""
from __future__ import annotations

import dataclasses
import numpy as np

FIXED_FIELDS = ("public", "gender", "completion", "age", "region", "clubs", "friends")
NUM_FIXED = 7


@dataclasses.dataclass
class World:
    text_columns: list
    uid: np.ndarray
    pub: np.ndarray
    comp: np.ndarray
    gender: np.ndarray
    age: np.ndarray
    region: np.ndarray
    clubs_ptr: np.ndarray
    clubs: np.ndarray
    friends_ptr: np.ndarray
    friends: np.ndarray
    tok_ptr: np.ndarray
    tok_id: np.ndarray
    tok_tf: np.ndarray
    adj_uid: np.ndarray
    adj_ptr: np.ndarray
    adj: np.ndarray
    norm_mean: np.ndarray = None
    norm_sd: np.ndarray = None
    norm_has: np.ndarray = None
    meta: dict = dataclasses.field(default_factory=dict)

    _DTYPES = dict(uid=np.int32, pub=np.int32, comp=np.int32, gender=np.int32, age=np.int32,
                   region=np.int32, clubs_ptr=np.int64, clubs=np.uint32, friends_ptr=np.int64,
                   friends=np.uint32, tok_ptr=np.int64, tok_id=np.int32, tok_tf=np.int32,
                   adj_uid=np.int32, adj_ptr=np.int64, adj=np.int32,
                   norm_mean=np.float32, norm_sd=np.float32, norm_has=np.uint8)

    def __post_init__(self):
        nf = NUM_FIXED + len(self.text_columns)
        if self.norm_mean is None:
            self.norm_mean = np.zeros(nf, np.float32)
            self.norm_sd = np.zeros(nf, np.float32)
            self.norm_has = np.zeros(nf, np.uint8)
        for k, dt in self._DTYPES.items():
            setattr(self, k, np.ascontiguousarray(getattr(self, k), dtype=dt))
        self.validate()

    # ------------------------------------------------------------------
    @property
    def n(self) -> int:
        return int(self.uid.shape[0])

    @property
    def T(self) -> int:
        return len(self.text_columns)

    @property
    def n_fields(self) -> int:
        return NUM_FIXED + self.T

    @property
    def n_adj_rows(self) -> int:
        return int(self.adj_uid.shape[0])

    def validate(self):
        n, T = self.n, self.T
        assert self.region.shape == (n, 3)
        for a in (self.pub, self.comp, self.gender, self.age):
            assert a.shape == (n,)
        assert self.clubs_ptr.shape == (n + 1,) and self.clubs_ptr[-1] == self.clubs.shape[0]
        assert self.friends_ptr.shape == (n + 1,) and self.friends_ptr[-1] == self.friends.shape[0]
        assert self.tok_ptr.shape == (n * T + 1,) and self.tok_ptr[-1] == self.tok_id.shape[0]
        assert self.tok_tf.shape == self.tok_id.shape
        assert self.adj_ptr.shape == (self.n_adj_rows + 1,) and self.adj_ptr[-1] == self.adj.shape[0]
        assert self.norm_mean.shape == (self.n_fields,)
        assert len(np.unique(self.uid)) == n, "duplicate profile uid"
        assert len(np.unique(self.adj_uid)) == self.n_adj_rows, "duplicate adjacency row"

    def set_normalizers(self, mean, sd, has=None):
        self.norm_mean = np.ascontiguousarray(mean, np.float32)
        self.norm_sd = np.ascontiguousarray(sd, np.float32)
        self.norm_has = (np.ones(self.n_fields, np.uint8) if has is None
                         else np.ascontiguousarray(has, np.uint8))
        assert self.norm_mean.shape == (self.n_fields,)

    # ------------------------------------------------------------------
    def canonical_record_bytes(self) -> np.ndarray:
        """canonical (encoding-independent) bytes of each profile record.
        16 (packed scalars + region) + 12 (three row offsets) + 4|clubs| + 4|friends|
        + sum over filled columns of (8 + 8 nnz)."""
        n, T = self.n, self.T
        cell = np.diff(self.tok_ptr).reshape(n, T)
        filled = (cell > 0).sum(1)
        nnz = cell.sum(1)
        return (28 + 4 * np.diff(self.clubs_ptr) + 4 * np.diff(self.friends_ptr)
                + 8 * filled + 8 * nnz).astype(np.int64)

    def head(self, m: int) -> "World":
        """The first `m` profiles (world order) with their adjacency rows; neighbours may point at users
        outside the slice (they then have no profile, which every implementation skips)."""
        m = int(min(m, self.n))
        T = self.T
        keep_rows = np.isin(self.adj_uid, self.uid[:m])
        idx = np.flatnonzero(keep_rows)
        lens = np.diff(self.adj_ptr)[idx]
        adj_ptr = np.zeros(len(idx) + 1, np.int64)
        np.cumsum(lens, out=adj_ptr[1:])
        if len(idx):
            take = np.concatenate([np.arange(self.adj_ptr[i], self.adj_ptr[i + 1]) for i in idx]) if len(idx) < 64 else \
                np.repeat(self.adj_ptr[idx] - adj_ptr[:-1], lens) + np.arange(adj_ptr[-1])
        else:
            take = np.zeros(0, np.int64)
        w = World(text_columns=list(self.text_columns), uid=self.uid[:m], pub=self.pub[:m], comp=self.comp[:m],
                  gender=self.gender[:m], age=self.age[:m], region=self.region[:m],
                  clubs_ptr=self.clubs_ptr[:m + 1], clubs=self.clubs[:self.clubs_ptr[m]],
                  friends_ptr=self.friends_ptr[:m + 1], friends=self.friends[:self.friends_ptr[m]],
                  tok_ptr=self.tok_ptr[:m * T + 1], tok_id=self.tok_id[:self.tok_ptr[m * T]],
                  tok_tf=self.tok_tf[:self.tok_ptr[m * T]], adj_uid=self.adj_uid[idx], adj_ptr=adj_ptr,
                  adj=self.adj[take], norm_mean=self.norm_mean, norm_sd=self.norm_sd, norm_has=self.norm_has)
        w.meta = dict(self.meta, head_of=self.n, n=m)
        return w

    def save(self, path):
        d = {k: getattr(self, k) for k in self._DTYPES}
        d["text_columns"] = np.array(self.text_columns)
        np.savez(path, **d)

    @classmethod
    def load(cls, path):
        z = np.load(path, allow_pickle=False)
        kw = {k: z[k] for k in cls._DTYPES}
        return cls(text_columns=[str(s) for s in z["text_columns"]], **kw)

    def subset_adj_lookup(self):
        """uid -> adjacency row index (python dict; tests / small worlds only)."""
        return {int(u): i for i, u in enumerate(self.adj_uid)}

    def uid_index(self):
        return {int(u): i for i, u in enumerate(self.uid)}
