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
