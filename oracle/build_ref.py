"""oracle/build_ref.py -- TEST INFRASTRUCTURE.

Compiles the UNMODIFIED reference (sources read where they lie under
/root/reference, nothing copied) into oracle/_ref/:

  openbnsllib<ext>     the reference's own pybind11 module (bindings/pybind_module.cpp)
  obn_refharness<ext>  oracle/ref_harness.cpp + the reference headers (OpenMP batch loop)

The reference's CMake build needs Boost (CMakeLists.txt:31), which this image
lacks; we therefore run g++ directly on its few source files with the two-file
Boost shim in oracle/ref_shim/ on the include path.  src/base/all_dims_cache.cpp
is skipped (dead code, and it does not compile: std::iota without <numeric>).

/root/reference does not exist on the GPU box; there the prebuilt .so files
(git-ignored, but shipped by gpurun) are used as they are.
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("OBN_REFERENCE_DIR", "/root/reference")
OUT = os.path.join(HERE, "_ref")
EXT = sysconfig.get_config_var("EXT_SUFFIX")


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources if os.path.exists(s))


def available():
    return os.path.exists(os.path.join(OUT, "openbnsllib" + EXT)) and os.path.exists(
        os.path.join(OUT, "obn_refharness" + EXT))


def build(verbose=False):
    """Build both modules if /root/reference is present. Returns True if usable."""
    if not os.path.isdir(os.path.join(REF, "include")):
        return available()
    os.makedirs(OUT, exist_ok=True)
    inc = [
        "-I", os.path.join(HERE, "ref_shim"),
        "-I", os.path.join(REF, "include"),
        "-I", os.path.join(REF, "external/pybind11/include"),
        "-I", sysconfig.get_paths()["include"],
        "-I", os.path.join(REF, "bindings"),
    ]
    flags = ["g++", "-std=c++17", "-O3", "-fopenmp", "-fPIC", "-shared",
             "-fvisibility=hidden", "-w"]
    shim = [os.path.join(HERE, "gammainc.h"),
            os.path.join(HERE, "ref_shim/boost/math/distributions/chi_squared.hpp")]
    jobs = []
    mod = os.path.join(OUT, "openbnsllib" + EXT)
    mod_src = [os.path.join(REF, p) for p in (
        "src/base/dataframe_wrapper.cpp",
        "src/graph/pdag_with_adjmat.cpp",
        "src/structure_learning/exhaustive_search.cpp",
        "src/structure_learning/pc.cpp",
        "src/structure_learning/rai.cpp",
        "src/structure_learning/simulated_annealing.cpp",
        "src/template.cpp",
        "bindings/pybind_module.cpp",
    )]
    if not _newer(mod, mod_src + shim):
        jobs.append(flags + inc + mod_src + ["-o", mod])
    har = os.path.join(OUT, "obn_refharness" + EXT)
    har_src = [os.path.join(HERE, "ref_harness.cpp"), os.path.join(REF, "src/graph/pdag_with_adjmat.cpp")]
    if not _newer(har, har_src + shim):
        jobs.append(flags + inc + har_src + ["-o", har])
    procs = [subprocess.Popen(j, stdout=subprocess.PIPE, stderr=subprocess.STDOUT) for j in jobs]
    ok = True
    for p, j in zip(procs, jobs):
        out, _ = p.communicate()
        if p.returncode != 0:
            ok = False
            sys.stderr.write("reference build failed: %s\n%s\n" % (" ".join(j), out.decode()))
        elif verbose:
            print("built", j[-1])
    return ok and available()


if __name__ == "__main__":
    print("reference oracle available:", build(verbose=True))
