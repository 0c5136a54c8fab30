"""Builds the native pieces in-tree (so the .so files travel with the repo snapshot).

    python -m dmlc_core_b200.build            # everything
    python -m dmlc_core_b200.build cuda host  # selected targets

Targets
  cuda    dmlc_core_b200/lib/libdmlc_b200.so      the C ABI + sm_100a kernels (nvcc)
  checked dmlc_core_b200/lib/libdmlc_b200_checked.so  same, bounds-checked (-DB200P_CHECKED, tests)
  host    dmlc_core_b200/lib/libdmlc_b200_host.so the dmlc::Parser<IndexType> C++ layer (g++)
  gen     dmlc_core_b200/lib/libb200gen.so        synthetic-corpus generators (gcc)
  emul    tests/_build/libstrtonum_emul.so        host build of strtonum_device.h (test only)
  oracle  oracle/_build/liboracle.so (+ oracle/_ref when /root/reference is present)
  driver  tests/_build/libsvm_parser_driver (the reference-style C++ throughput driver)
"""
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "dmlc_core_b200")
LIB = os.path.join(PKG, "lib")
NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd):
    sys.stderr.write("[build] " + " ".join(cmd) + "\n")
    env = dict(os.environ)
    env.pop("CXX", None)
    env.pop("CC", None)
    subprocess.check_call(cmd, cwd=ROOT, env=env)


def _glob(d, exts):
    out = []
    for base, _, files in os.walk(d):
        for f in files:
            if f.endswith(exts):
                out.append(os.path.join(base, f))
    return out


def build_cuda(force=False):
    os.makedirs(LIB, exist_ok=True)
    target = os.path.join(LIB, "libdmlc_b200.so")
    srcs = _glob(os.path.join(PKG, "csrc"), (".cu", ".cuh", ".h")) + \
        [os.path.join(ROOT, "include", "dmlc_b200_parser.h")]
    if force or _stale(target, srcs):
        _run([_nvcc()] + NVCC_ARCH + ["-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
                                      "-shared", "--cudart", "static", "-o", target,
                                      os.path.join(PKG, "csrc", "b200_parser.cu")])
    return target


def build_cuda_checked(force=False):
    """The same library with -DB200P_CHECKED: bounds-checked queues and stores (test build)."""
    os.makedirs(LIB, exist_ok=True)
    target = os.path.join(LIB, "libdmlc_b200_checked.so")
    srcs = _glob(os.path.join(PKG, "csrc"), (".cu", ".cuh", ".h")) + \
        [os.path.join(ROOT, "include", "dmlc_b200_parser.h")]
    if force or _stale(target, srcs):
        _run([_nvcc()] + NVCC_ARCH + ["-lineinfo", "-O3", "-std=c++17", "-DB200P_CHECKED", "-Xcompiler",
                                      "-fPIC", "-shared", "--cudart", "static", "-o", target,
                                      os.path.join(PKG, "csrc", "b200_parser.cu")])
    return target


def build_host(force=False):
    os.makedirs(LIB, exist_ok=True)
    target = os.path.join(LIB, "libdmlc_b200_host.so")
    src_dir = os.path.join(PKG, "src")
    if not os.path.isdir(src_dir):
        return None
    srcs = _glob(src_dir, (".cc", ".h")) + _glob(os.path.join(PKG, "include"), (".h",))
    ccs = [s for s in srcs if s.endswith(".cc")]
    if not ccs:
        return None
    if force or _stale(target, srcs + [os.path.join(ROOT, "include", "dmlc_b200_parser.h"),
                                       os.path.join(PKG, "csrc", "host_pool.h")]):
        _run(["g++", "-O2", "-std=c++14", "-fPIC", "-shared", "-pthread", "-Wall",
              "-I" + os.path.join(PKG, "include"), "-I" + os.path.join(ROOT, "include"),
              "-o", target] + ccs + ["-L" + LIB, "-ldmlc_b200", "-Wl,-rpath,$ORIGIN", "-ldl"])
    return target


def build_gen(force=False):
    os.makedirs(LIB, exist_ok=True)
    target = os.path.join(LIB, "libb200gen.so")
    src = os.path.join(PKG, "csrc", "datagen.c")
    if not os.path.exists(src):
        return None
    if force or _stale(target, [src]):
        _run(["gcc", "-O2", "-std=gnu11", "-fPIC", "-shared", "-fopenmp", "-o", target, src, "-lm"])
    return target


def build_emul(force=False):
    out_dir = os.path.join(ROOT, "tests", "_build")
    os.makedirs(out_dir, exist_ok=True)
    target = os.path.join(out_dir, "libstrtonum_emul.so")
    src = os.path.join(ROOT, "tests", "strtonum_emul.cc")
    if not os.path.exists(src):
        return None
    if force or _stale(target, [src, os.path.join(PKG, "csrc", "strtonum_device.h")]):
        _run(["g++", "-O2", "-msse2", "-ffp-contract=off", "-std=c++14", "-fPIC", "-shared",
              "-DB200P_HOST_EMULATION", "-I" + os.path.join(PKG, "csrc"), "-o", target, src])
    return target


def build_driver(force=False):
    """tests/cpp/libsvm_parser_driver.cc against this package's headers (and, where /root/reference
    exists, the same source against the reference's headers + oracle/_ref for comparison)."""
    out_dir = os.path.join(ROOT, "tests", "_build")
    os.makedirs(out_dir, exist_ok=True)
    src = os.path.join(ROOT, "tests", "cpp", "libsvm_parser_driver.cc")
    target = os.path.join(out_dir, "libsvm_parser_driver")
    host = os.path.join(LIB, "libdmlc_b200_host.so")
    if os.path.exists(host) and (force or _stale(target, [src, host])):
        _run(["g++", "-std=c++14", "-O2", "-I" + os.path.join(PKG, "include"), src, "-L" + LIB,
              "-ldmlc_b200_host", "-ldmlc_b200", "-Wl,-rpath,$ORIGIN/../../dmlc_core_b200/lib",
              "-o", target])
    ref = os.path.join(ROOT, "oracle", "_ref", "libdmlc_ref.so")
    target_ref = os.path.join(out_dir, "libsvm_parser_driver_ref")
    if os.path.isdir("/root/reference/include") and os.path.exists(ref) and \
            (force or _stale(target_ref, [src, ref])):
        _run(["g++", "-std=c++11", "-O2", "-fopenmp", "-I/root/reference/include", src, ref,
              "-Wl,-rpath,$ORIGIN/../../oracle/_ref", "-o", target_ref])
    return target


def build_oracle(force=False):
    odir = os.path.join(ROOT, "oracle")
    _run(["make", "-s", "-C", odir, "oracle"])
    ref = os.path.join(odir, "_ref", "libdmlc_ref.so")
    if os.path.isdir("/root/reference/src") and (force or _stale(
            ref, [os.path.join(odir, "ref_shim.cc"), os.path.join(odir, "Makefile")])):
        _run(["make", "-s", "-C", odir, "ref"])
    dropin = os.path.join(odir, "_ref", "libdmlc_dropin.so")
    if os.path.isdir("/root/reference/src") and os.path.exists(os.path.join(LIB, "libdmlc_b200.so")) \
            and (force or _stale(dropin, _glob(os.path.join(PKG, "src"), (".cc", ".h")) +
                                 [os.path.join(odir, "Makefile")])):
        _run(["make", "-s", "-C", odir, "dropin"])
    return os.path.join(odir, "_build", "liboracle.so")


TARGETS = {"cuda": build_cuda, "host": build_host, "gen": build_gen, "emul": build_emul,
           "oracle": build_oracle, "driver": build_driver, "checked": build_cuda_checked}


def build_all(force=False):
    for name in ("cuda", "checked", "gen", "host", "emul", "oracle", "driver"):
        TARGETS[name](force)


if __name__ == "__main__":
    names = [a for a in sys.argv[1:] if not a.startswith("-")]
    force = "--force" in sys.argv
    if not names:
        build_all(force)
    else:
        for n in names:
            TARGETS[n](force)
