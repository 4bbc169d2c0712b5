"""Writes a `World` in the reference's on-disk formats, so that the reference's own
binaries (api_cli, kurs) can load it:

  <dir>/config/text_columns.txt      one column name per line             (src/utils.cpp:13-24)
  <dir>/data/users_encoded.csv       user_id,public,completion_percentage,gender,region,age,clubs,friends,<col>_tokens...
                                     (src/encoder.cpp:171-173, parsed positionally by src/user_loader.cpp:20-29)
  <dir>/data/adjacency.csv           uid,n1,n2,...                        (src/graph_builder.cpp:39-75)
  <dir>/data/column_normalizers.csv  column,mean,stddev                   (src/utils.cpp:123-153)
  <dir>/data/tokens.csv, clubs_map.csv  headers (+ club names)            (src/vocab_builder.cpp:133-197)

The loader reads at most 100 000 rows (src/user_loader.cpp:34)."""
from __future__ import annotations

import os
import numpy as np

from .world import World, FIXED_FIELDS


def write_reference_data_dir(w: World, root: str, club_names: bool = True) -> str:
    os.makedirs(os.path.join(root, "config"), exist_ok=True)
    data = os.path.join(root, "data")
    os.makedirs(data, exist_ok=True)
    with open(os.path.join(root, "config", "text_columns.txt"), "w") as f:
        f.write("\n".join(w.text_columns) + "\n")
    T = w.T
    with open(os.path.join(data, "users_encoded.csv"), "w") as f:
        f.write("user_id,public,completion_percentage,gender,region,age,clubs,friends," +
                ",".join(c + "_tokens" for c in w.text_columns) + "\n")
        for i in range(w.n):
            reg = ";".join("" if int(x) < 0 else str(int(x)) for x in w.region[i])
            clubs = ";".join(str(int(x)) for x in w.clubs[w.clubs_ptr[i]:w.clubs_ptr[i + 1]])
            friends = ";".join(str(int(x)) for x in w.friends[w.friends_ptr[i]:w.friends_ptr[i + 1]])
            cells = []
            for t in range(T):
                b, e = w.tok_ptr[i * T + t], w.tok_ptr[i * T + t + 1]
                cells.append(";".join(f"{int(a)}:{int(c)}" for a, c in zip(w.tok_id[b:e], w.tok_tf[b:e])))
            pub = "" if int(w.pub[i]) < 0 else str(int(w.pub[i]))
            gen = "" if int(w.gender[i]) < 0 else str(int(w.gender[i]))
            f.write(f"{int(w.uid[i])},{pub},{int(w.comp[i])},{gen},{reg},{int(w.age[i])},{clubs},{friends}," + ",".join(cells) + "\n")
    with open(os.path.join(data, "adjacency.csv"), "w") as f:
        for r in np.argsort(w.adj_uid, kind="stable"):
            row = w.adj[w.adj_ptr[r]:w.adj_ptr[r + 1]]
            f.write(",".join([str(int(w.adj_uid[r]))] + [str(int(x)) for x in row]) + "\n")
    names = list(FIXED_FIELDS) + list(w.text_columns)
    with open(os.path.join(data, "column_normalizers.csv"), "w") as f:
        f.write("column,mean,stddev\n")
        for k, name in enumerate(names):
            if w.norm_has[k]:
                f.write(f"{name},{float(w.norm_mean[k]):.9g},{float(w.norm_sd[k]):.9g}\n")
    with open(os.path.join(data, "tokens.csv"), "w") as f:
        f.write("column,token,tid,df\n")
    with open(os.path.join(data, "clubs_map.csv"), "w") as f:
        f.write("club_id,slug,title\n")
        if club_names:
            for cid in np.unique(w.clubs):
                f.write(f"{int(cid)},club-{int(cid)},Club {int(cid)}\n")
    return root
