#!/bin/bash
timeout 40 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "snapshot_parity or chunked_transfer or streaming" 2>&1 | tail -2
