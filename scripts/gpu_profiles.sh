#!/bin/bash
# ncu evidence for profiles/: (1) launch list of one decode token (graph off, PDL on), (2) --set full of the dominant decode
# GEMV kernels and the attention kernel, (3) --set full of the tcgen05 prefill GEMM, (4) launch list of one batched step.
set -u
mkdir -p gpurun_out
cat > /tmp/decode_probe.py <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
from aios_b200 import capi, gguf_synth as gs
import bench
path = bench.ensure_model("mistral-7b", "Q4_K_M", True)
eng = capi.Engine(0, 1, capi.ENGINE_NO_GRAPH | capi.ENGINE_EXACT_PREFILL)
m = eng.load_model("m", path, n_ctx=512)
s = m.new_seq()
s.prefill(gs.synthetic_prompt("mistral-7b", 384, bench.SEED))
s.decode(4)
import ctypes
rt = ctypes.CDLL("libcudart.so.12")
rt.cudaProfilerStart()
s.decode(2)
rt.cudaProfilerStop()
print("ok")
PY
cat > /tmp/tc_probe.py <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
from aios_b200 import capi, gguf_synth as gs
import bench
path = bench.ensure_model("mistral-7b", "Q8_0", True)
eng = capi.Engine(0, 1, 0)
m = eng.load_model("m", path, n_ctx=4100)
s = m.new_seq()
os.environ["AIOS_PROFILE_STEPS"] = "1"
ms = s.bench_prefill(gs.synthetic_prompt("mistral-7b", 4096, 5), 1)
print("prefill ms", ms)
PY
python /tmp/decode_probe.py > gpurun_out/decode_probe_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_decode_launches.csv python /tmp/decode_probe.py > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'gemv_db|attn_split' -s 170 -c 8 -o gpurun_out/r02_decode_full python /tmp/decode_probe.py > gpurun_out/ncu2.log 2>&1
python /tmp/tc_probe.py > gpurun_out/tc_probe_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'prefill_tc|attn_prefill' -s 40 -c 6 -o gpurun_out/r02_tc_full python /tmp/tc_probe.py > gpurun_out/ncu3.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_prefill_launches.csv python /tmp/tc_probe.py > gpurun_out/ncu4.log 2>&1
python scripts/batch_probe.py mistral-7b 64 > gpurun_out/probe_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_batch_launches.csv python scripts/batch_probe.py mistral-7b 64 > gpurun_out/ncu5.log 2>&1
tail -2 gpurun_out/tc_probe_plain.log gpurun_out/probe_plain.log; ls -la gpurun_out | tail -12
