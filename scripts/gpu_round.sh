#!/bin/bash
# One gpurun call: GPU tests, smoke, bench (ours + reference arm), launch list, micro-probes.  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
./scripts/ubench/ubench > gpurun_out/ubench.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc $?" >> gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc $?" >> gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
tail -3 gpurun_out/pytest.log; cat gpurun_out/smoke.log | tail -2; cat gpurun_out/bench.json | head -c 1500
# launch list of the same short bench command (cold-cache, serialised: compare SHARES)
timeout 600 python bench.py --steps 1 --warmup 3 --gen 2 --no-cpu-baseline --no-subrecords > gpurun_out/ncu_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'gemv|attn|sample|embed|gemm|quant' -s 2000 -c 400 --csv \
    --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --gen 2 --no-cpu-baseline --no-subrecords > gpurun_out/ncu.log 2>&1
