#!/bin/bash
# Round-end evidence run on one B200: GPU test suite, smoke(), the bench line as the driver runs it and
# with default flags, and the ncu launch list of a short bench run (numbers under ncu are never bench values).
python -m pytest tests -m gpu -x -q > gpurun_out/r03_gpu_suite_final.log 2>&1; echo "suite rc=$?" >> gpurun_out/r03_gpu_suite_final.log
tail -3 gpurun_out/r03_gpu_suite_final.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r03_bench_n1.json 2> gpurun_out/r03_bench_n1.err; echo "bench rc=$?"
python bench.py > gpurun_out/r03_bench_n1_default.json 2> gpurun_out/r03_bench_default.err; echo "bench-default rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r03_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-sustained --no-e2e > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r03_bench_n1.json", "gpurun_out/r03_bench_n1_default.json"):
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d["steps"], round(d["value"]), round(d["ms_per_step"], 4), round(d["roofline"]["frac"], 4),
          round(d["e2e"]["value"]), d["parity_check"]["mismatch"], d["e2e"]["parity_check"]["mismatch"], d["clocks"]["reasons"])
PY
