#!/bin/bash
# usage: scripts/gpu_quick.sh "<pytest args>" -- run a subset with a timeout, log to gpurun_out/quick.log
mkdir -p gpurun_out
timeout 600 python -m pytest $1 -q -m gpu --timeout=120 > gpurun_out/quick.log 2>&1
echo "exit $?"; tail -40 gpurun_out/quick.log
