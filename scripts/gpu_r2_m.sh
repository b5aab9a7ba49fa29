#!/bin/bash
mkdir -p gpurun_out

python scripts/snap_e2e_probe.py 2>&1 | grep call

