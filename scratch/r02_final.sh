#!/bin/bash
# the round-end sequence of the driver, in one call: GPU tests, smoke, the default bench line
v=${1:-11}; b=${2:-10}
python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/r02_gpu_tests_v$v.log; cat gpurun_out/r02_gpu_tests_v$v.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r02_bench_v$b.json 2> gpurun_out/r02_bench_v$b.err; tail -c 600 gpurun_out/r02_bench_v$b.err
python - <<PY
import json
d = json.load(open("gpurun_out/r02_bench_v$b.json"))
print(d["value"], d["e2e"]["value"], d["e2e"]["compact"]["value"], d["roofline"]["kernel"], d["roofline"]["frac"])
print({k: (v.get("msamples_per_s") if isinstance(v, dict) else v) for k, v in d["extras"].items()})
PY
