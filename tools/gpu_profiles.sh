#!/bin/bash
# One GPU session that produces the judged evidence: parity tests, the bench line, the ncu launch list of the same
# bench command, and one `ncu --set full` capture of each hot kernel (GEMM, row quantiser).
# Usage (from the repo root, on the GPU box): bash tools/gpu_profiles.sh <tag>      e.g. r01
TAG=${1:-r01}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -2 gpurun_out/smoke.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -1 gpurun_out/${TAG}_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>> gpurun_out/bench.err
tail -1 gpurun_out/${TAG}_bench_reference.json
# launch list of the bench command (2 decoder layers so that the list stays short; same kernels, same shapes)
BCMD="python bench.py --steps 2 --warmup 3 --layers 2 --no-cpu-baseline"
$BCMD > gpurun_out/bench_l2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"mixed_gemm|quant_rows|colstat|argsort" -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $BCMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$? lines=$(wc -l < gpurun_out/${TAG}_launches.csv)"
# DRAM bytes per launch of the same command (roofline.traffic): one step of 2 layers = 2 x (7 GEMM + 4 quantiser)
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"mixed_gemm|quant_rows" -c 400 --csv \
    --log-file gpurun_out/${TAG}_traffic.csv $BCMD > gpurun_out/ncu_traffic.log 2>&1
python tools/traffic_summary.py gpurun_out/${TAG}_traffic.csv 22 gpurun_out/${TAG}_traffic.json
# full captures: one GEMM launch and one quantiser launch at the 7B prefill shape
GCMD="python tools/gemm_time.py 32768 4096 4096 128 0.05 2 0"
$GCMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:mixed_gemm -s 2 -c 1 -f -o gpurun_out/${TAG}_gemm $GCMD \
    > gpurun_out/ncu_gemm.log 2>&1
echo "ncu gemm rc=$?"
ncu --set full --clock-control none --import-source on -k regex:quant_rows -s 0 -c 1 -f -o gpurun_out/${TAG}_quant $GCMD \
    > gpurun_out/ncu_quant.log 2>&1
echo "ncu quant rc=$?"
ls -la gpurun_out/ | tail -20
