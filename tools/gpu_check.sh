#!/bin/bash
# usage: tools/gpu_check.sh <tag> [extra command]   — GPU parity tests + per-kernel timing of C2 on a B200 box
tag=$1; shift
/usr/local/graft/bin/gpurun --timeout 900 -- "timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; tail -4 gpurun_out/${tag}_tests.log; timeout 200 python tools/kt.py C2 30 2>&1 | tee gpurun_out/${tag}_kt.log; $*" 2>&1 | tail -25
