#!/bin/bash
# One piece of judged evidence per GPU call (one profiler run per call, after the same command ran clean without it):
#   bash tools/gpu_evidence.sh suite|launches|traffic|gemm|k1|k1norm|decode [tag]
MODE=$1; TAG=${2:-r02}
mkdir -p gpurun_out
BCMD="python bench.py --steps 2 --warmup 3 --layers 2 --no-cpu-baseline --no-gpu-baselines --no-decode"
case $MODE in
suite)
  python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log
  tail -4 gpurun_out/${TAG}_pytest_gpu.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${TAG}_smoke.log
  tail -2 gpurun_out/${TAG}_smoke.log
  python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
  python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>> gpurun_out/${TAG}_bench.err; echo "ref rc=$?"
  tail -c 1500 gpurun_out/${TAG}_bench_reference.json
  for a in "32768 4096 128 0.05" "32768 4096 128 0.05 1" "32768 11008 128 0.05" "32768 11008 128 0.05 1" "32768 8192 128 0.10"; do
    python tools/k1_time.py $a 2>&1 | grep K1T; done > gpurun_out/${TAG}_k1_time.log
  for n in rms layer unfused unfused_layer; do MQ_NORM=$n python tools/k1_time.py 32768 4096 128 0.05 2>&1 | grep K1T; done >> gpurun_out/${TAG}_k1_time.log
  cat gpurun_out/${TAG}_k1_time.log
  python tools/decode_time.py 1 16 64 2>&1 | grep DEC > gpurun_out/${TAG}_decode_time.log; tail -3 gpurun_out/${TAG}_decode_time.log
  ;;
launches)
  $BCMD > gpurun_out/bench_l2.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"mixed_gemm|quant_rows|colstat|argsort|decode_gemm" -c 400 --csv \
      --log-file gpurun_out/${TAG}_launches.csv $BCMD > gpurun_out/ncu_launches.log 2>&1
  echo "rc=$? lines=$(wc -l < gpurun_out/${TAG}_launches.csv)"
  python tools/launch_summary.py gpurun_out/${TAG}_launches.csv 22 | tee gpurun_out/${TAG}_launches_summary.txt
  ;;
traffic)
  $BCMD > gpurun_out/bench_l2.log 2>&1 && \
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"mixed_gemm|quant_rows" -c 400 --csv \
      --log-file gpurun_out/${TAG}_traffic.csv $BCMD > gpurun_out/ncu_traffic.log 2>&1
  echo "rc=$?"; python tools/traffic_summary.py gpurun_out/${TAG}_traffic.csv 22 gpurun_out/${TAG}_traffic.json
  ;;
gemm)
  C="python tools/gemm_time.py 32768 4096 4096 128 0.05 2 0"
  $C > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:mixed_gemm -s 2 -c 1 -f \
      -o gpurun_out/${TAG}_gemm $C > gpurun_out/ncu_gemm.log 2>&1; echo "rc=$?"
  python tools/ncu_details.py gpurun_out/${TAG}_gemm.ncu-rep gpurun_out/${TAG}_gemm_ncu_details.txt; grep -iE "tensor|Issue Slots|Duration|DRAM Throughput|Registers" gpurun_out/${TAG}_gemm_ncu_details.txt | head -30
  ;;
k1)
  C="python tools/k1_time.py 32768 4096 128 0.05"
  $C > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:quant_rows_h2 -s 3 -c 1 -f \
      -o gpurun_out/${TAG}_k1 $C > gpurun_out/ncu_k1.log 2>&1; echo "rc=$?"
  python tools/ncu_details.py gpurun_out/${TAG}_k1.ncu-rep gpurun_out/${TAG}_k1_ncu_details.txt; grep -iE "Issue Slots|Duration|DRAM Throughput|Memory Throughput|Registers" gpurun_out/${TAG}_k1_ncu_details.txt | head -20
  ;;
k1norm)
  C="python tools/k1_time.py 32768 4096 128 0.05"
  MQ_NORM=rms $C > gpurun_out/plain.log 2>&1 && MQ_NORM=rms ncu --set full --clock-control none --import-source on -k regex:quant_rows_h2 -s 3 -c 1 -f \
      -o gpurun_out/${TAG}_k1norm $C > gpurun_out/ncu_k1norm.log 2>&1; echo "rc=$?"
  python tools/ncu_details.py gpurun_out/${TAG}_k1norm.ncu-rep gpurun_out/${TAG}_k1norm_ncu_details.txt; grep -iE "Issue Slots|Duration|DRAM Throughput|Memory Throughput|Registers" gpurun_out/${TAG}_k1norm_ncu_details.txt | head -20
  ;;
decode)
  C="python tools/decode_one.py 16 28672 8192 5"
  $C > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:decode_gemm_kernel -s 3 -c 1 -f \
      -o gpurun_out/${TAG}_decode $C > gpurun_out/ncu_decode.log 2>&1; echo "rc=$?"
  python tools/ncu_details.py gpurun_out/${TAG}_decode.ncu-rep gpurun_out/${TAG}_decode_ncu_details.txt; grep -iE "Issue Slots|Duration|DRAM Throughput|Memory Throughput|Registers" gpurun_out/${TAG}_decode_ncu_details.txt | head -20
  ;;
esac
ls -la gpurun_out | tail -8
