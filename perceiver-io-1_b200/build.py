"""Build libpio_b200.so (all CUDA kernels + the C ABI) in-tree with nvcc for sm_100a.

    python perceiver-io-1_b200/build.py [--force]

The .so lands next to this file so it travels with the repo snapshot to the GPU box.
"""
import concurrent.futures
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libpio_b200.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "-I", INCLUDE]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, extra_flags=(), lib=None, obj_dir=None):
    """extra_flags / lib / obj_dir: instrumented variants (e.g. -DPIO_LB_TRACE -> libpio_b200_trace.so for tools/latent_trace.py)"""
    global OBJ, LIB
    if lib is not None:
        LIB = lib
    if obj_dir is not None:
        OBJ = obj_dir
    os.makedirs(OBJ, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(INCLUDE, "*.h"))
    jobs = []
    for s in srcs:
        o = os.path.join(OBJ, os.path.basename(s)[:-3] + ".o")
        if force or _stale(o, [s] + hdrs):
            jobs.append((s, o))

    def compile_one(so):
        s, o = so
        cmd = [NVCC] + FLAGS + list(extra_flags) + ["-c", s, "-o", o]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        return s, r.returncode, r.stdout + r.stderr

    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for s, rc, out in ex.map(compile_one, jobs):
            if out.strip() and (verbose or rc != 0):
                print(out, file=sys.stderr)
            if rc != 0:
                raise RuntimeError(f"nvcc failed on {s}")
    objs = [os.path.join(OBJ, os.path.basename(s)[:-3] + ".o") for s in srcs]
    if force or jobs or _stale(LIB, objs):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            print(r.stdout + r.stderr, file=sys.stderr)
            raise RuntimeError("link failed")
    return LIB


if __name__ == "__main__":
    if "--trace" in sys.argv:
        print(build(force=True, extra_flags=["-DPIO_LB_TRACE"], lib=os.path.join(HERE, "libpio_b200_trace.so"),
                    obj_dir=os.path.join(HERE, "build_trace")))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
