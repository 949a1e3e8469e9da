#!/bin/bash
# GEMM iteration loop: correctness probe, then timings (with the wait-cycle profile) at the 7B prefill shapes
mkdir -p gpurun_out
timeout 150 python tools/gemm_probe.py > gpurun_out/gemm_probe.log 2>&1
grep -v '"nbad": 0' gpurun_out/gemm_probe.log | head -5; echo "probe ok cases: $(grep -c '"nbad": 0' gpurun_out/gemm_probe.log) of $(grep -c CASE gpurun_out/gemm_probe.log)"
export MQ_PROFILE=${MQ_PROFILE-1}
{
for g in 128 1024 64; do timeout 120 python tools/gemm_time.py 32768 4096 4096 $g 0.05 10 0; done
timeout 120 python tools/gemm_time.py 32768 11008 4096 128 0.05 10 0
timeout 120 python tools/gemm_time.py 32768 4096 11008 128 0.05 10 0
timeout 120 python tools/gemm_time.py 32768 4096 4096 128 0.0 10 0
} > gpurun_out/gemm_prof.log 2>&1
grep -h "GEMM\|PROF\|Error\|error" gpurun_out/gemm_prof.log | cut -c1-330
