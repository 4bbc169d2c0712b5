"""Builds dropin/_build/libpokec_dropin.so: the reference's UNCHANGED include/recommender.h and
include/tfidf_index.h implemented by dropin/recommender_b200.cpp + dropin/tfidf_index_b200.cpp on top of
libpokec_b200.so, wrapped by the same C driver the compiled-reference checker uses (oracle/ref_driver.cpp,
-DPK_DROPIN_BUILD).  The result exports the same ref_* entry points as oracle/_ref/libpokec_ref.so, so a
test can put the CUDA-backed `Recommender` and the reference's CPU `Recommender` side by side, both driven
through the reference's own C++ API.

Needs the reference headers, so it only builds where /root/reference exists (this container); the .so is
git-ignored and travels to the GPU box with the gpurun snapshot.  No reference source of the REPLACED path is compiled
in; src/test.cpp (the hold-out harness, a caller of recommend_collaborative) is, unmodified, from where it lies.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("POKEC_REFERENCE_DIR", "/root/reference")
OUT_DIR = os.path.join(HERE, "_build")
OUT = os.path.join(OUT_DIR, "libpokec_dropin.so")
PKG = os.path.join(ROOT, "recommendation_system_pokec_b200")


def build(force=False):
    if not os.path.isdir(os.path.join(REF, "include")):
        return OUT if os.path.exists(OUT) else None
    from recommendation_system_pokec_b200.build import build as build_cuda
    build_cuda()
    srcs = [os.path.join(HERE, "recommender_b200.cpp"), os.path.join(HERE, "tfidf_index_b200.cpp"),
            os.path.join(ROOT, "oracle", "ref_driver.cpp"),
            os.path.join(REF, "src", "test.cpp")]      # the reference's hold-out harness: caller code, compiled where it lies
    deps = srcs + [os.path.join(ROOT, "include", "pokec_b200.h"), os.path.join(ROOT, "include", "pokec_world.h"),
                   os.path.join(HERE, "recommender_b200_batch.h")]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(d) <= os.path.getmtime(OUT) for d in deps):
        return OUT
    os.makedirs(OUT_DIR, exist_ok=True)
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-pthread", "-DPK_DROPIN_BUILD",
           "-I", os.path.join(REF, "include"), "-I", os.path.join(ROOT, "include"), "-I", HERE] + srcs + \
          ["-L", PKG, "-lpokec_b200", "-Wl,-rpath,$ORIGIN/../../recommendation_system_pokec_b200", "-o", OUT]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    sys.path.insert(0, ROOT)
    print(build(force="--force" in sys.argv))
