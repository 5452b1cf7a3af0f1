#!/bin/bash
# parity suite + smoke + the default bench line
set -u
mkdir -p gpurun_out
timeout 1000 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
tail -4 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1]); print(round(d["value"]), round(d["ms_per_step"],3), d["e2e"], d["roofline"]["frac"], d["config"]["pipeline"])
PY
