#!/bin/bash
# r02 GPU session A: full parity suite (all BASELINE configs), smoke, bench line.
mkdir -p gpurun_out; rm -f gpurun_out/parity_log.jsonl
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/a_smi.txt 2>&1
nproc >> gpurun_out/a_smi.txt
timeout 2400 python -m pytest tests -m gpu -q --durations=20 > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -40 gpurun_out/a_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/a_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/a_smoke.log
tail -2 gpurun_out/a_smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/a_bench.json; tail -5 gpurun_out/a_bench.err
