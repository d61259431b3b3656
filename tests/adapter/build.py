"""TEST INFRASTRUCTURE: builds the mjpc::Planner-interface harness around mujoco_mpc_b200/host/mjpc_adapter (needs the
reference headers, i.e. /root/reference; the built libraries travel to the GPU box with the snapshot).
  libadapter_host.so : adapter + host planner class + the host stand-in of the engine ABI (tests/hostsim/hostengine.cc)
  libadapter_cuda.so : adapter + host planner class, linked against mujoco_mpc_b200/libps_engine.so"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "..", "..")
REF = os.environ.get("REF", "/root/reference")


def _stale(out, deps):
    return not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps if os.path.exists(d))


def build(kind):
    """kind: 'host' or 'cuda'.  Returns the library path, or None when it neither exists nor can be built here."""
    out = os.path.join(HERE, f"libadapter_{kind}.so")
    if not os.path.isdir(os.path.join(REF, "mjpc")):
        return out if os.path.exists(out) else None
    host = os.path.join(ROOT, "mujoco_mpc_b200", "host")
    refso = os.path.join(ROOT, "oracle", "_ref")
    srcs = [os.path.join(HERE, "adapter_harness.cc"), os.path.join(host, "mjpc_adapter", "cuda_sampling_adapter.cc"),
            os.path.join(host, "cuda_sampling_planner.cc")]
    deps = srcs + [os.path.join(host, f) for f in os.listdir(host) if f.endswith(".h")] + \
        [os.path.join(host, "mjpc_adapter", f) for f in os.listdir(os.path.join(host, "mjpc_adapter"))] + [os.path.join(refso, "libmjpc_ref.so")]
    cmd = ["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-I" + os.path.join(ROOT, "oracle", "ref_shim"), "-I" + REF, "-o", out]
    if kind == "host":
        srcs.append(os.path.join(HERE, "..", "hostsim", "hostengine.cc"))
        csrc = os.path.join(ROOT, "mujoco_mpc_b200", "csrc")
        deps += [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cuh", ".h"))]
        cmd += ["-Wl,-Bsymbolic"]
    if not _stale(out, deps):
        return out
    for s in srcs:
        cmd += ["-x", "c++", s]
    cmd += ["-x", "none", "-L" + refso, "-lmjpc_ref", "-Wl,-rpath," + os.path.abspath(refso)]
    if kind == "cuda":
        eng = os.path.join(ROOT, "mujoco_mpc_b200")
        cmd += ["-L" + eng, "-lps_engine", "-Wl,-rpath," + os.path.abspath(eng)]
    subprocess.run(cmd + ["-lpthread"], check=True)
    return out
