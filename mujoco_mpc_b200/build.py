"""In-tree build of the CUDA engine: nvcc, sm_100a only, one translation unit per engine variant
compiled in parallel, linked into mujoco_mpc_b200/libps_engine.so (the C-ABI library)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libps_engine.so")
UNITS = ["ps_engine.cu", "ps_variant_tiny.cu", "ps_variant_planar.cu", "ps_variant_legged.cu",
         "ps_variant_tiny_coop.cu", "ps_variant_planar_coop.cu", "ps_variant_legged_coop.cu", "ps_variant_planar_sparse.cu",
         "ps_variant_legged_sparse.cu", "ps_variant_robust.cu", "ps_noalloc.cu", "ps_variant_tiny_spec.cu", "ps_variant_legged_xl.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
# -fmad=false: FP64 multiply-adds are NOT contracted, so the device arithmetic is operation-for-operation
# the IEEE sequence the oracle executes (-ffp-contract=off).  Contact-rich rollouts amplify rounding
# differences by many orders of magnitude, so this is what keeps the 1e-9 / 1e-8 parity bars robust.
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]
if os.environ.get("PS_FMAD", "0") != "1":
    FLAGS.append("-fmad=false")


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, extra=()):
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs += [os.path.join(HERE, "..", "include", f) for f in ("ps_engine.h", "ps_model.h")]
    bdir = os.path.join(HERE, "build")
    os.makedirs(bdir, exist_ok=True)
    objs, jobs = [], []
    for u in UNITS:
        src = os.path.join(CSRC, u)
        obj = os.path.join(bdir, u.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + hdrs):
            cmd = [NVCC] + FLAGS + list(extra) + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or r.returncode:
            sys.stderr.write(r.stdout + r.stderr)
        if r.returncode:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))

    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(run, jobs))
    if jobs or force or _stale(OUT, objs):
        run([NVCC, "-shared", "-o", OUT] + objs + ["-lcudart"])
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))


def build_host(force=False):
    """Host-side planner library (C++ mirror of mjpc::SamplingPlanner above the C ABI)."""
    hdir = os.path.join(HERE, "host")
    if not os.path.isdir(hdir):
        return None
    srcs = [os.path.join(hdir, f) for f in sorted(os.listdir(hdir)) if f.endswith(".cc")]
    hdrs = [os.path.join(hdir, f) for f in os.listdir(hdir) if f.endswith(".h")]
    out = os.path.join(HERE, "libps_planner.so")
    if srcs and (force or _stale(out, srcs + hdrs)):
        cmd = [os.environ.get("CXX", "g++"), "-std=c++17", "-O2", "-fPIC", "-shared", "-o", out] + srcs + \
              ["-L" + HERE, "-lps_engine", "-Wl,-rpath,$ORIGIN", "-lpthread"]
        subprocess.check_call(cmd)
    return out
