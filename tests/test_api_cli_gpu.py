"""End-to-end drop-in check: the reference's own server binary `api_cli`, built once from the
unmodified reference (oracle/_ref/api_cli_ref) and once with its five scoring translation units replaced by
dropin/*.cpp + libpokec_b200.so (dropin/_build/api_cli_b200) -- api_cli.cpp itself unmodified in both -- is run
on the same data directory; every JSON line the two print after READY must be identical byte for byte
(scores are printed with 6 decimals, src/api_cli.cpp:217)."""
import os
import subprocess

import numpy as np
import pytest

from recommendation_system_pokec_b200.refdata import write_reference_data_dir
from recommendation_system_pokec_b200.synth import sample_queries
from tests.conftest import make_test_world

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "api_cli_ref")
OUR_BIN = os.path.join(ROOT, "dropin", "_build", "api_cli_b200")


def run(binary, cwd, text):
    r = subprocess.run([binary], input=text, capture_output=True, text=True, cwd=cwd, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = r.stdout.splitlines()
    assert "READY" in lines, r.stdout[:500]
    return lines[lines.index("READY") + 1:], r.stderr


def test_api_cli_json_identical(tmp_path):
    if not (os.path.exists(REF_BIN) and os.path.exists(OUR_BIN)):
        pytest.skip("api_cli binaries not built (dropin/build_api_cli.py needs the reference sources)")
    w = make_test_world(n=1500, seed=23, pairs=800)
    w.age[::7] = 0                                     # exercised by fill_missing_ages (src/user_loader.cpp:131-140)
    d = str(tmp_path)
    write_reference_data_dir(w, d)
    users = [int(u) for u in sample_queries(w, 40, seed=6)]
    text = "PING\n" + "".join(f"USER {u}\n" for u in users) + "USER 987654321\n\nBOGUS\nEXIT\n"
    ref_lines, _ = run(REF_BIN, d, text)
    our_lines, our_err = run(OUR_BIN, d, text)
    assert len(ref_lines) == len(our_lines) == len(users) + 5
    bad = [i for i, (a, b) in enumerate(zip(ref_lines, our_lines)) if a != b]
    assert not bad, f"{len(bad)} of {len(ref_lines)} lines differ; first: line {bad[0]}\nref: {ref_lines[bad[0]][:600]}\nour: {our_lines[bad[0]][:600]}"
    assert sum('"recommendations"' in ln for ln in our_lines) == len(users)
    assert "[pokec_b200] fatal" not in our_err


BATCH_BIN = os.path.join(ROOT, "dropin", "_build", "api_cli_b200_batch")


def test_api_cli_batch_lines_equal_user_lines(tmp_path):
    """the BATCH line (dropin/api_cli_batch.inc, one device call for many users) returns, for every
    user and each of the four modes, exactly the list the REFERENCE binary prints in its USER reply; the USER /
    PING / error lines of the patched server stay byte-identical to the reference's."""
    import json
    if not (os.path.exists(REF_BIN) and os.path.exists(BATCH_BIN)):
        pytest.skip("api_cli binaries not built (dropin/build_api_cli.py needs the reference sources)")
    w = make_test_world(n=1500, seed=29, pairs=800)
    d = str(tmp_path)
    write_reference_data_dir(w, d)
    users = [int(u) for u in sample_queries(w, 70, seed=9)]          # 70 users: three query groups in one call
    user_lines = "".join(f"USER {u}\n" for u in users)
    ref_lines, _ = run(REF_BIN, d, "PING\n" + user_lines + "BOGUS\nEXIT\n")
    ulist = " ".join(str(u) for u in users)
    modes = ["graph", "collaborative", "interest", "clubs"]
    text = "PING\n" + user_lines + "BOGUS\n" + "".join(f"BATCH {m} 20 {ulist}\n" for m in modes) + \
           f"BATCH graph 20 {users[0]} 987654321 {users[1]}\nBATCH graph 7 {users[2]}\n" \
           "BATCH nosuchmode 20 1\nBATCH graph 0 1\nBATCH graph 20\nBATCH graph 20 12 x\nEXIT\n"
    our_lines, our_err = run(BATCH_BIN, d, text)
    assert "[pokec_b200] fatal" not in our_err
    n0 = len(users) + 2
    assert our_lines[:n0] == ref_lines[:n0], "USER / PING / unknown-command lines changed under the BATCH patch"
    ref_recs = [json.loads(ln)["recommendations"] for ln in ref_lines[1:1 + len(users)]]
    for mi, m in enumerate(modes):
        rep = json.loads(our_lines[n0 + mi])
        assert rep["mode"] == m and rep["k"] == 20 and len(rep["results"]) == len(users)
        for u, res, rr in zip(users, rep["results"], ref_recs):
            assert res["user_id"] == u
            assert res["items"] == rr[m], f"BATCH {m} differs from the reference's USER reply for user {u}"
    mixed = json.loads(our_lines[n0 + 4])["results"]
    assert mixed[0]["items"] == ref_recs[0]["graph"] and mixed[2]["items"] == ref_recs[1]["graph"]
    assert mixed[1] == {"user_id": 987654321, "error": "not found"}
    k7 = json.loads(our_lines[n0 + 5])
    assert k7["k"] == 7 and k7["results"][0]["items"] == ref_recs[2]["graph"][:7]
    assert our_lines[n0 + 6:n0 + 10] == ['{"error":"bad batch"}'] * 4
    assert json.loads(our_lines[n0 + 10]) == {"ok": True, "exiting": True}


def test_api_cli_save_and_batch_fullscan(tmp_path):
    """SURVEY 8 f1 + f2 through the server: SAVE writes the world the server actually holds (CSV loaders,
    fill_missing_ages and the parsed normalizers applied) as a world file; `BATCH fullscan` -- the north-star path,
    FAS top-k against every profile -- must then equal the CPU oracle run on exactly that file."""
    import json
    from oracle.oracle import CpuOracle
    from recommendation_system_pokec_b200 import engine
    from recommendation_system_pokec_b200.cabi import EXCLUDE_SELF_AND_ADJ
    if not os.path.exists(BATCH_BIN):
        pytest.skip("api_cli_b200_batch not built (dropin/build_api_cli.py needs the reference sources)")
    w = make_test_world(n=1500, seed=31, pairs=800)
    w.age[::9] = 0
    d = str(tmp_path)
    write_reference_data_dir(w, d)
    users = [int(u) for u in sample_queries(w, 40, seed=12)]
    path = os.path.join(d, "world.pkw")
    text = f"SAVE {path}\nBATCH fullscan 50 {' '.join(map(str, users))}\nSAVE\nBATCH fullscan 257 {users[0]}\nEXIT\n"
    lines, err = run(BATCH_BIN, d, text)
    assert "[pokec_b200] fatal" not in err
    assert json.loads(lines[0]) == {"ok": True, "saved": path}
    assert lines[2] == '{"error":"save failed"}' and lines[3] == '{"error":"bad batch"}'
    with engine.WorldFile(path) as f:
        served = f.to_world(w.text_columns)
    assert served.n == w.n and not (served.age <= 0).any()          # the server filled the missing ages
    orc = CpuOracle(served, kind="port")
    want_id, want_sc, want_cnt = orc.fullscan(np.array(users, np.int32), 50, EXCLUDE_SELF_AND_ADJ, nthreads=8)
    rep = json.loads(lines[1])
    assert rep["mode"] == "fullscan" and rep["k"] == 50 and len(rep["results"]) == len(users)
    for i, (u, res) in enumerate(zip(users, rep["results"])):
        assert res["user_id"] == u and len(res["items"]) == int(want_cnt[i])
        assert [it["id"] for it in res["items"]] == [int(x) for x in want_id[i, :want_cnt[i]]], f"user {u}"
        got = np.array([it["score"] for it in res["items"]])
        assert np.abs(got - want_sc[i, :want_cnt[i]].astype(np.float64)).max() <= 2e-6
    orc.close()
