"""Build the test-infrastructure libraries (never part of the product path):

  oracle/libpokec_oracle.so      CPU restatement (oracle/fas_oracle.cpp), always built
  oracle/_ref/libpokec_ref.so    the UNMODIFIED reference compiled from its own sources where they
                                 lie under /root/reference (+ oracle/ref_driver.cpp, a C ABI around
                                 the reference's public Recommender API).  Built only where
                                 /root/reference exists (this container); the prebuilt .so travels to
                                 the GPU box with the gpurun snapshot (oracle/_ref/ is git-ignored).

The reference's own build system (cmake) is not used: the path compiles from six of its source files.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("POKEC_REFERENCE_DIR", "/root/reference")
REF_SOURCES = ["recommender.cpp", "recommender_similarity.cpp", "recommender_graph.cpp",
               "recommender_clubs.cpp", "tfidf_index.cpp", "utils.cpp",
               "test.cpp"]      # test.cpp = the hold-out harness, a CALLER of the path (ref_holdout_test)


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build_restatement(force=False):
    src = os.path.join(HERE, "fas_oracle.cpp")
    out = os.path.join(HERE, "libpokec_oracle.so")
    deps = [src, os.path.join(ROOT, "include", "pokec_world.h")]
    if force or _newer(out, deps):
        cmd = ["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-pthread",
               "-I", os.path.join(ROOT, "include"), src, "-o", out]
        subprocess.check_call(cmd)
    return out


def build_reference(force=False):
    """-O3, the optimisation level of the reference's Release build; no -march flags (baseline x86-64)."""
    out_dir = os.path.join(HERE, "_ref")
    out = os.path.join(out_dir, "libpokec_ref.so")
    if not os.path.isdir(os.path.join(REF, "src")):
        return out if os.path.exists(out) else None
    os.makedirs(out_dir, exist_ok=True)
    srcs = [os.path.join(REF, "src", f) for f in REF_SOURCES]
    drv = os.path.join(HERE, "ref_driver.cpp")
    deps = srcs + [drv, os.path.join(ROOT, "include", "pokec_world.h")]
    if force or _newer(out, deps):
        cmd = ["g++", "-O3", "-std=c++17", "-shared", "-fPIC", "-pthread", "-w",
               "-I", os.path.join(REF, "include"), "-I", os.path.join(ROOT, "include"),
               drv] + srcs + ["-o", out]
        subprocess.check_call(cmd)
    return out


def build_all(force=False):
    return build_restatement(force), build_reference(force)


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv))
