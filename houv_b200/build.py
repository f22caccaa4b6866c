"""Ahead-of-time build of libhouv_b200.so (the C-ABI library) with nvcc for sm_100a.

No torch headers are involved: the library is plain CUDA runtime code behind the
C ABI in include/houv_b200.h.  The .so is built IN-TREE (houv_b200/libhouv_b200.so)
so that it travels with a repo snapshot; the reference JIT-compiles at import
instead (utils/metrics/CD/chamfer3D/dist_chamfer_3D.py:12-16).

No --use_fast_math: FTZ / approximate division would break bit parity of idx.
"""
import glob
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libhouv_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O2",
    "--fmad=false",  # only explicit fma intrinsics fuse: arithmetic is spelled out
]


def _nvcc():
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + _headers()
    return any(os.path.getmtime(d) > t for d in deps)


def _headers():
    return glob.glob(os.path.join(CSRC, "*.cuh")) + [
        os.path.join(HERE, "..", "include", "houv_b200.h"), os.path.abspath(__file__)]


def build(force=False, verbose=False, extra=()):
    """Compile every .cu under csrc/ into one shared library. Returns its path.

    Safe under torchrun: every rank may call this at import.  The whole build runs under an
    exclusive file lock (the first rank compiles, the others wait and then find it up to date),
    objects are rebuilt only when older than their source or any header, and the library is
    linked to a temporary name and os.replace()d into place, so no process can ever dlopen a
    half-written file."""
    if not force and not needs_build():
        return LIB
    import fcntl
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    with open(os.path.join(objdir, ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not needs_build():  # another process built it while we waited
                return LIB
            nvcc = _nvcc()
            hdr_t = max(os.path.getmtime(h) for h in _headers())
            objs, procs = [], []
            for s in sources():
                o = os.path.join(objdir, os.path.basename(s)[:-3] + ".o")
                objs.append(o)
                if not force and not extra and os.path.exists(o) and \
                        os.path.getmtime(o) > max(hdr_t, os.path.getmtime(s)):
                    continue
                cmd = [nvcc] + NVCC_FLAGS + list(extra) + (["-Xptxas", "-v"] if verbose else []) + [
                    "-c", s, "-o", o]
                procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE,
                                                    stderr=subprocess.STDOUT)))
            for cmd, p in procs:
                out = p.communicate()[0].decode()
                if verbose or p.returncode:
                    sys.stderr.write(out)
                if p.returncode:
                    raise RuntimeError("nvcc failed: " + " ".join(cmd))
            tmp = LIB + ".tmp.%d" % os.getpid()
            subprocess.check_call([nvcc, "-shared", "-o", tmp] + objs + ["-lcudart"])
            os.replace(tmp, LIB)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
