#!/bin/bash
mkdir -p gpurun_out
nvcc -std=c++17 -gencode arch=compute_100a,code=sm_100a -o /tmp/gather4_probe scripts/experiments/gather4_probe.cu > gpurun_out/r2t_build.log 2>&1; echo "build rc=$?"
timeout 60 /tmp/gather4_probe > gpurun_out/r2t_gather4_probe.txt 2>&1; echo "probe rc=$?"; cat gpurun_out/r2t_gather4_probe.txt
