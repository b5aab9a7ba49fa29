#!/bin/bash
timeout 120 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q -k "delta_pk or pk_bin" 2>&1 | tail -3
