#!/bin/bash
# usage: tools/gpu_quick.sh <pytest args...>
mkdir -p gpurun_out
timeout 900 python -m pytest "$@" -q -m gpu -x --timeout 600 2>&1 | tail -30
