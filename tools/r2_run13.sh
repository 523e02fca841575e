#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_run13_tests.log 2>&1; echo "tests exit $?"; tail -6 gpurun_out/r2_run13_tests.log
