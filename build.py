#!/usr/bin/env python
"""In-tree build of the native pieces (no JIT cache, no site-packages install).

* ``metisfl_b200/lib/libmetisfl_b200_kernels.so`` — sm_100a CUDA kernels behind the C ABI of
  ``include/metisfl_b200.h`` (nvcc, ``-gencode arch=compute_100a,code=sm_100a -lineinfo``).
* ``metisfl_b200/_core*.so`` — host C++17 controller core (pybind11, g++ -O3 -fopenmp).

Usage: ``python build.py [kernels|core|all] [--force] [-v]``.
"""
import glob
import hashlib
import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
LIBDIR = os.path.join(ROOT, "metisfl_b200", "lib")
OBJDIR = os.path.join(ROOT, "build")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-I", os.path.join(ROOT, "include"),
              "-I", os.path.join(ROOT, "csrc")]


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("build step failed: " + " ".join(cmd))
    if verbose and (r.stdout or r.stderr):
        print(r.stdout + r.stderr)


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _headers():
    return (glob.glob(os.path.join(ROOT, "include", "*.h")) +
            glob.glob(os.path.join(ROOT, "csrc", "**", "*.cuh"), recursive=True) +
            glob.glob(os.path.join(ROOT, "csrc", "**", "*.h"), recursive=True) +
            glob.glob(os.path.join(ROOT, "csrc", "**", "*.hpp"), recursive=True))


def build_kernels(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(ROOT, "csrc", "kernels", "*.cu")) +
                  glob.glob(os.path.join(ROOT, "csrc", "ckks", "*.cu")))
    hdrs = _headers()
    objs, jobs = [], []
    for s in srcs:
        o = os.path.join(OBJDIR, os.path.basename(s).replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            extra = ["-Xptxas", "-v"] if verbose else []
            jobs.append([NVCC] + NVCC_FLAGS + extra + ["-c", s, "-o", o])
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(lambda c: _run(c, verbose), jobs))
    out = os.path.join(LIBDIR, "libmetisfl_b200_kernels.so")
    if jobs or force or _stale(out, objs):
        _run([NVCC, "-shared", "-o", out] + objs + ["-cudart", "static", "-gencode",
             "arch=compute_100a,code=sm_100a"], verbose)
    return out


def build_core(force=False, verbose=False):
    import pybind11
    srcs = sorted(glob.glob(os.path.join(ROOT, "csrc", "core", "*.cc")) +
                  glob.glob(os.path.join(ROOT, "csrc", "ckks", "*.cc")))
    if not srcs:
        return None
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    out = os.path.join(ROOT, "metisfl_b200", "_core" + ext)
    if force or _stale(out, srcs + _headers()):
        # -ffp-contract=off: the aggregation rules round every w * x term before adding it (reference semantics, golden
        # vectors); a fused multiply-add in the AVX-512 clones of the inner loops would skip that rounding
        cmd = ["g++", "-O3", "-std=c++17", "-shared", "-fPIC", "-fopenmp", "-ffp-contract=off", "-fvisibility=hidden",
               "-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"],
               "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "csrc")] + srcs + ["-o", out]
        _run(cmd, verbose)
    return out


def build_tools(force=False, verbose=False):
    """Standalone native binaries (aggregation micro-benchmark of the controller core)."""
    srcs = sorted(glob.glob(os.path.join(ROOT, "csrc", "tools", "*.cc")))
    core = [s for s in sorted(glob.glob(os.path.join(ROOT, "csrc", "core", "*.cc")) +
                              glob.glob(os.path.join(ROOT, "csrc", "ckks", "*.cc"))) if not s.endswith("pybind.cc")]
    outs = []
    os.makedirs(os.path.join(ROOT, "metisfl_b200", "bin"), exist_ok=True)
    for s in srcs:
        out = os.path.join(ROOT, "metisfl_b200", "bin", os.path.basename(s)[:-3].replace("_main", ""))
        if force or _stale(out, [s] + core + _headers()):
            _run(["g++", "-O3", "-std=c++17", "-fopenmp", "-ffp-contract=off", "-pthread", "-I", os.path.join(ROOT, "include"),
                  "-I", os.path.join(ROOT, "csrc"), s] + core + ["-o", out], verbose)
        outs.append(out)
    return outs


def main(argv):
    what = [a for a in argv if not a.startswith("-")] or ["all"]
    force, verbose = "--force" in argv, "-v" in argv
    if "kernels" in what or "all" in what:
        print("built", build_kernels(force, verbose))
    if "core" in what or "all" in what:
        print("built", build_core(force, verbose))
    if "tools" in what or "all" in what:
        print("built", build_tools(force, verbose))


if __name__ == "__main__":
    main(sys.argv[1:])
