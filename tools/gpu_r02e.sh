#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
export LIFEB200_TRACE_GATHER=1
timeout 300 tests/cpp/_build/multi_survey 16777216 4096 0 nocheck 2>&1 | tee gpurun_out/r02e_cpp_trace_n$N.txt
timeout 300 tests/cpp/_build/multi_survey 16777216 64 0 nocheck 2>&1 | tee -a gpurun_out/r02e_cpp_trace_n$N.txt
