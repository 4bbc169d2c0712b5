"""ctypes view of include/pokec_world.h (struct pk_world) built from a `World`."""
from __future__ import annotations

import ctypes as C
import numpy as np

from .world import World

MODE_GRAPH, MODE_COLLAB, MODE_CLUBS = 0, 1, 2
EXCLUDE_NONE, EXCLUDE_SELF, EXCLUDE_SELF_AND_ADJ = 0, 1, 2

i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)
f32p = C.POINTER(C.c_float)


class PkWorld(C.Structure):
    _fields_ = [
        ("n_users", C.c_int32), ("n_text_cols", C.c_int32),
        ("uid", i32p), ("pub", i32p), ("comp", i32p), ("gender", i32p), ("age", i32p), ("region", i32p),
        ("clubs_ptr", i64p), ("clubs", u32p), ("friends_ptr", i64p), ("friends", u32p),
        ("tok_ptr", i64p), ("tok_id", i32p), ("tok_tf", i32p),
        ("n_adj_rows", C.c_int32), ("adj_uid", i32p), ("adj_ptr", i64p), ("adj", i32p),
        ("norm_mean", f32p), ("norm_sd", f32p), ("norm_has", u8p),
    ]


def ptr(a: np.ndarray, ty):
    return a.ctypes.data_as(ty)


def make_pk_world(w: World) -> PkWorld:
    """The returned struct borrows the World's arrays: keep `w` alive while it is in use."""
    s = PkWorld()
    s.n_users, s.n_text_cols, s.n_adj_rows = w.n, w.T, w.n_adj_rows
    for name, ty in (("uid", i32p), ("pub", i32p), ("comp", i32p), ("gender", i32p), ("age", i32p),
                     ("region", i32p), ("clubs_ptr", i64p), ("clubs", u32p), ("friends_ptr", i64p),
                     ("friends", u32p), ("tok_ptr", i64p), ("tok_id", i32p), ("tok_tf", i32p),
                     ("adj_uid", i32p), ("adj_ptr", i64p), ("adj", i32p), ("norm_mean", f32p),
                     ("norm_sd", f32p), ("norm_has", u8p)):
        setattr(s, name, ptr(getattr(w, name), ty))
    s._keepalive = w
    return s


def col_names_array(w: World):
    arr = (C.c_char_p * max(w.T, 1))()
    for i, name in enumerate(w.text_columns):
        arr[i] = name.encode()
    return arr
