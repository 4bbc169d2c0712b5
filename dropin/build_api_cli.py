"""Builds the reference's server binary twice, with g++ directly from the sources where they lie under
/root/reference (not the reference's cmake):

  oracle/_ref/api_cli_ref       every reference translation unit (the CPU reference as shipped)
  dropin/_build/api_cli_b200    the same, with src/recommender.cpp, recommender_similarity.cpp, recommender_graph.cpp,
                                recommender_clubs.cpp and tfidf_index.cpp REPLACED by dropin/*.cpp + libpokec_b200.so

  dropin/_build/api_cli_b200_batch  api_cli_b200 plus the BATCH line: src/api_cli.cpp is read where it
                                lies, one #include and the handler dropin/api_cli_batch.inc are inserted at two
                                anchors, and the patched text is written to dropin/_build/ (a build product,
                                git-ignored) -- the patch a maintainer would make by hand, INTEGRATION.md section 6

api_cli.cpp and every other caller compile unmodified against the drop-in -- that is the integration claim
(INTEGRATION.md); tests/test_api_cli_gpu.py runs both on the same data directory and compares their JSON lines.
All outputs are git-ignored and travel to the GPU box with the gpurun snapshot."""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("POKEC_REFERENCE_DIR", "/root/reference")
REPLACED = {"recommender.cpp", "recommender_similarity.cpp", "recommender_graph.cpp", "recommender_clubs.cpp", "tfidf_index.cpp"}
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "api_cli_ref")
OUR_BIN = os.path.join(HERE, "_build", "api_cli_b200")
BATCH_BIN = os.path.join(HERE, "_build", "api_cli_b200_batch")
PKG = os.path.join(ROOT, "recommendation_system_pokec_b200")
INCLUDE_ANCHOR = '#include "recommender.h"'
HANDLER_ANCHOR = 'if (cmd == "USER" && uid >= 0) {'


def _patched_api_cli():
    """src/api_cli.cpp with the BATCH handler: returns the path of the patched copy under dropin/_build/"""
    with open(os.path.join(REF, "src", "api_cli.cpp")) as f:
        text = f.read()
    with open(os.path.join(HERE, "api_cli_batch.inc")) as f:
        handler = f.read()
    if text.count(INCLUDE_ANCHOR) != 1 or text.count(HANDLER_ANCHOR) != 1:
        raise RuntimeError("api_cli.cpp changed: the BATCH patch anchors were not found exactly once")
    text = text.replace(INCLUDE_ANCHOR, INCLUDE_ANCHOR + '\n#include "recommender_b200_batch.h"', 1)
    i = text.index(HANDLER_ANCHOR)
    i = text.rfind("\n", 0, i) + 1                    # start of the anchor's line
    text = text[:i] + handler + text[i:]
    out = os.path.join(HERE, "_build", "api_cli_batch.cpp")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    if not os.path.exists(out) or open(out).read() != text:
        with open(out, "w") as f:
            f.write(text)
    return out


def _objects(out_dir, srcs, extra):
    os.makedirs(out_dir, exist_ok=True)
    objs = []
    for s in srcs:
        o = os.path.join(out_dir, os.path.basename(s) + ".o")
        if not os.path.exists(o) or os.path.getmtime(o) < os.path.getmtime(s):
            subprocess.check_call(["g++", "-O3", "-std=c++17", "-w", "-c", s, "-o", o,
                                   "-I", os.path.join(REF, "include"), "-I", os.path.join(REF, "config"),
                                   "-I", os.path.join(REF, "third_party", "lemmagen", "include")] + extra)
        objs.append(o)
    return objs


def build(force=False):
    if not os.path.isdir(os.path.join(REF, "src")):
        return (REF_BIN if os.path.exists(REF_BIN) else None, OUR_BIN if os.path.exists(OUR_BIN) else None)
    sys.path.insert(0, ROOT)
    from recommendation_system_pokec_b200.build import build as build_cuda
    build_cuda()
    all_src = sorted(glob.glob(os.path.join(REF, "src", "*.cpp")) + glob.glob(os.path.join(REF, "third_party", "lemmagen", "src", "*.cpp")))
    core = [s for s in all_src if os.path.basename(s) not in ("main.cpp",)]
    common = [s for s in core if os.path.basename(s) not in REPLACED]
    ref_only = [s for s in core if os.path.basename(s) in REPLACED]
    obj_dir = os.path.join(ROOT, "oracle", "_ref", "obj")
    if force:
        for f in glob.glob(os.path.join(obj_dir, "*.o")) + glob.glob(os.path.join(HERE, "_build", "obj", "*.o")):
            os.remove(f)
    o_common = _objects(obj_dir, common, [])
    o_ref = _objects(obj_dir, ref_only, [])
    subprocess.check_call(["g++", "-O3", "-pthread"] + o_common + o_ref + ["-o", REF_BIN])
    ours = [os.path.join(HERE, "recommender_b200.cpp"), os.path.join(HERE, "tfidf_index_b200.cpp")]
    o_ours = _objects(os.path.join(HERE, "_build", "obj"), ours, ["-I", os.path.join(ROOT, "include"), "-I", HERE])
    subprocess.check_call(["g++", "-O3", "-pthread"] + o_common + o_ours +
                          ["-L", PKG, "-lpokec_b200", "-Wl,-rpath,$ORIGIN/../../recommendation_system_pokec_b200", "-o", OUR_BIN])
    # the same server with the BATCH line: api_cli.cpp patched at build time, everything else shared
    o_rest = [o for o in o_common if os.path.basename(o) != "api_cli.cpp.o"]
    o_batch = _objects(os.path.join(HERE, "_build", "obj"), [_patched_api_cli()], ["-I", HERE, "-I", os.path.join(ROOT, "include")])
    o_ours = _objects(os.path.join(HERE, "_build", "obj"), ours, ["-I", os.path.join(ROOT, "include"), "-I", HERE])
    subprocess.check_call(["g++", "-O3", "-pthread"] + o_rest + o_batch + o_ours +
                          ["-L", PKG, "-lpokec_b200", "-Wl,-rpath,$ORIGIN/../../recommendation_system_pokec_b200", "-o", BATCH_BIN])
    return REF_BIN, OUR_BIN


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
