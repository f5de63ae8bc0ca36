#!/bin/bash
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/r3e_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r3e_pytest.log
