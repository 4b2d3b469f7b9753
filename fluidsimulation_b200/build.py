"""Builds libfluidsim3_b200.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libfluidsim3_b200.so")
SOURCES = ["particles.cu", "levelset.cu", "p2g.cu", "project.cu", "shard.cu", "fs3_api.cu", "fs3_hooks.cu", "fs3_state.cu", "consumer.cu", "levelset_lines.cu", "no_alloc.cu"]
# -fmad=false: the reference is an MSVC /fp:precise build without FMA contraction; parity needs the same.
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-fmad=false",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-O2"]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, extra_flags=(), suffix=""):
    """suffix/extra_flags: instrumented side builds (tools/), e.g. suffix="_trace", extra_flags=["-DFS3_LS_TRACE"]."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    headers = [os.path.join(CSRC, "fs3_internal.cuh"), os.path.join(os.path.dirname(HERE), "include", "fs3.h")]
    objdir = os.path.join(HERE, "build" + suffix)
    lib = LIB.replace(".so", suffix + ".so")
    os.makedirs(objdir, exist_ok=True)
    objs, procs = [], []
    for src in SOURCES:
        sp = os.path.join(CSRC, src)
        op = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(op)
        if force or _stale(op, [sp] + headers):
            cmd = [nvcc] + NVCC_FLAGS + list(extra_flags) + (["-Xptxas", "-v"] if verbose else []) + ["-c", sp, "-o", op]
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError("nvcc failed on " + src)
    if force or procs or _stale(lib, objs):
        subprocess.check_call([nvcc, "-shared", "-o", lib] + objs + ["-lcudart"])
    return lib


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
