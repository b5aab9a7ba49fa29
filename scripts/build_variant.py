#!/usr/bin/env python
"""Link a tuning variant of the library: one source recompiled with extra -D flags, every other object as built.

    python scripts/build_variant.py snap_pf0 bf_snapshot.cu -DSN_PREFETCH=0
    BF_LIB_PATH=baryonification_b200/variants/libbf_snap_pf0.so python bench.py --workload snapshot ...
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baryonification_b200 import build as b  # noqa: E402

name, src, flags = sys.argv[1], sys.argv[2], sys.argv[3:]
b.build()
vdir = os.path.join(b.HERE, "variants")
os.makedirs(vdir, exist_ok=True)
obj = os.path.join(vdir, name + "_" + src[:-3] + ".o")
base = [f for f in b.NVCC_FLAGS if f != "--shared"]
subprocess.run(["nvcc"] + base + flags + ["-c", "-o", obj, os.path.join(b.CSRC, src)], check=True, capture_output=True)
objs = [os.path.join(b.HERE, "build", os.path.basename(s)[:-3] + ".o") for s in b.sources() if os.path.basename(s) != src]
so = os.path.join(vdir, "libbf_" + name + ".so")
subprocess.run(["nvcc", "--shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", so, obj] + objs, check=True)
print(so)
