#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -s -k "full_width or labels_large" 2>&1 | grep -v "^$" | tail -12
timeout 1500 python scripts/onf_sweep.py --json gpurun_out/onf_sweep.json > gpurun_out/onf_sweep.txt 2>&1; echo sweep rc=$?; tail -12 gpurun_out/onf_sweep.txt
