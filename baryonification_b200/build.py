"""Build the CUDA extension in-tree: baryonification_b200/libbaryonification_b200.so (sm_100a only)."""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libbaryonification_b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-fmad=false",  # index arithmetic must round like the CPU reference; hot loops call fma() explicitly
              "-Xcompiler", "-fPIC", "-Xcompiler", "-O2", "--shared", "-Xptxas", "-v"]


def _flag_key():
    return " ".join([f for f in NVCC_FLAGS if f != "--shared"] + os.environ.get("BF_NVCC_EXTRA", "").split())


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def needs_build():
    if not os.path.exists(SO):
        return True
    stamp = os.path.join(HERE, "build", "flags.txt")
    if os.path.exists(stamp) and open(stamp).read() != _flag_key():  # built with other flags (BF_NVCC_EXTRA)
        return True
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        [os.path.join(os.path.dirname(HERE), "include", "baryonification_b200.h")]
    return any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps)


def _run(cmd, env):
    res = subprocess.run(cmd, env=env, capture_output=True, text=True)
    return res.returncode, " ".join(cmd) + "\n" + res.stdout + res.stderr


def build(force=False, verbose=False):
    """Each .cu is compiled to an object (only when it or a header changed, in parallel), then linked."""
    if not (force or needs_build()):
        return SO
    from concurrent.futures import ThreadPoolExecutor
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    env = dict(os.environ)
    env.pop("CC", None), env.pop("CXX", None)  # use the system gcc as nvcc's host compiler
    extra = os.environ.get("BF_NVCC_EXTRA", "").split()
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(os.path.dirname(HERE), "include", "baryonification_b200.h")]
    hdr_time = max(os.path.getmtime(h) for h in hdrs)
    flags = [f for f in NVCC_FLAGS if f != "--shared"]
    stamp = os.path.join(objdir, "flags.txt")
    flag_key = _flag_key()
    if not os.path.exists(stamp) or open(stamp).read() != flag_key:
        force = True
    jobs, objs = [], []
    for src in sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_time):
            jobs.append([nvcc] + flags + extra + ["-c", "-o", obj, src])
    logs = []
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        results = list(ex.map(lambda c: _run(c, env), jobs))
    rc = max([r[0] for r in results] + [0])
    logs += [r[1] for r in results]
    if rc == 0:
        r = _run([nvcc, "--shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", SO] + objs, env)
        rc, logs = r[0], logs + [r[1]]
    log = os.path.join(HERE, "build.log")
    with open(log, "w") as f:
        f.write("\n".join(logs))
    if rc != 0:
        sys.stderr.write("\n".join(logs))
        raise RuntimeError("nvcc failed; see " + log)
    with open(stamp, "w") as f:
        f.write(flag_key)
    if verbose:
        print("\n".join(logs))
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
