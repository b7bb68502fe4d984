#!/usr/bin/env python
"""Build every native component in-tree (no JIT cache, no site-packages install).

  cxxnet_b200/lib/libcxxnet_b200.so   sm_100a kernel library (csrc/kernels/*.cu)
  cxxnet_b200/lib/libcxxnet_io.so     native data pipeline (csrc/runtime/*.cpp)
  cxxnet_b200/lib/libcxxnetwrapper.so the 22-function C ABI (csrc/capi)
  bin/im2bin                          image list packer (tools/im2bin.cpp)

Usage: python build.py [--force] [--verbose]
"""
import glob
import os
import shutil
import subprocess
import sys
import sysconfig

ROOT = os.path.dirname(os.path.abspath(__file__))
LIBDIR = os.path.join(ROOT, "cxxnet_b200", "lib")
NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--use_fast_math", "-Xptxas", "-v"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd, verbose, log=None):
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log:
        with open(log, "w") as f:
            f.write(r.stdout)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("build step failed: %s" % " ".join(cmd))
    elif verbose:
        print(r.stdout)


def build_kernels(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(ROOT, "build", "kernels")
    os.makedirs(objdir, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(ROOT, "csrc", "kernels", "*.cu")))
    hdrs = glob.glob(os.path.join(ROOT, "csrc", "kernels", "*.cuh")) + glob.glob(os.path.join(ROOT, "include", "*.h"))
    objs = []
    procs = []
    for s in srcs:
        o = os.path.join(objdir, os.path.basename(s)[:-3] + ".o")
        objs.append(o)
        if force or _newer(o, [s] + hdrs):
            cmd = [NVCC] + ARCH + NVCC_FLAGS + ["-I", os.path.join(ROOT, "include"), "-c", s, "-o", o]
            if verbose:
                print(" ".join(cmd), flush=True)
            procs.append((s, o, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for s, o, p in procs:
        out, _ = p.communicate()
        with open(o + ".log", "w") as f:
            f.write(out)
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError("nvcc failed for %s" % s)
    target = os.path.join(LIBDIR, "libcxxnet_b200.so")
    if force or _newer(target, objs):
        _run([NVCC] + ARCH + ["-shared", "-o", target] + objs + ["-lcudart_static", "-ldl", "-lrt", "-lpthread"], verbose)
    return target


def build_host(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    cxx = os.environ.get("CXX", "g++")
    flags = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-pthread", "-ffp-contract=off"]
    out = []
    io_srcs = sorted(glob.glob(os.path.join(ROOT, "csrc", "runtime", "*.cpp")))
    io_hdrs = glob.glob(os.path.join(ROOT, "csrc", "runtime", "*.h"))
    if io_srcs:
        target = os.path.join(LIBDIR, "libcxxnet_io.so")
        if force or _newer(target, io_srcs + io_hdrs):
            _run([cxx] + flags + ["-O3", "-shared", "-o", target] + io_srcs + ["-lz"], verbose)
        out.append(target)
    capi = sorted(glob.glob(os.path.join(ROOT, "csrc", "capi", "*.cpp")))
    if capi:
        target = os.path.join(LIBDIR, "libcxxnetwrapper.so")
        inc = sysconfig.get_paths()["include"]
        libdir = sysconfig.get_config_var("LIBDIR")
        ver = sysconfig.get_config_var("LDVERSION") or sysconfig.get_python_version()
        if force or _newer(target, capi + glob.glob(os.path.join(ROOT, "include", "*.h"))):
            _run([cxx] + flags + ["-shared", "-o", target, "-I", inc, "-I", os.path.join(ROOT, "include")] + capi +
                 ["-L", libdir, "-lpython" + ver, "-Wl,-rpath," + libdir, "-ldl"], verbose)
        out.append(target)
    tool = os.path.join(ROOT, "tools", "im2bin.cpp")
    if os.path.exists(tool):
        os.makedirs(os.path.join(ROOT, "bin"), exist_ok=True)
        target = os.path.join(ROOT, "bin", "im2bin")
        if force or _newer(target, [tool] + io_hdrs + io_srcs):
            _run([cxx] + flags + ["-o", target, "-I", os.path.join(ROOT, "csrc", "runtime"), tool, "-lz"], verbose)
        out.append(target)
    return out


def build_all(force=False, verbose=False):
    k = build_kernels(force, verbose)
    h = build_host(force, verbose)
    return [k] + h


if __name__ == "__main__":
    outs = build_all(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    for o in outs:
        print("built", os.path.relpath(o, ROOT))
