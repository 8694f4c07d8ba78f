#!/bin/bash
# usage: scripts/quick_train.sh <tag>  -- parity tests of the CNN path + the train-only bench, one GPU
tag=$1
timeout 600 python -m pytest tests/test_bench_sizes_gpu.py tests/test_net_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 400 python bench.py --steps 30 --warmup 5 --no-preprocess --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo rc=$?
python - <<PY
import json
j=json.loads(open("gpurun_out/${tag}_bench.json").read().strip().splitlines()[-1])
print(j["ms_per_step"], j["train_b256"]["ms_per_step"], j["train_c5_deepnet"]["ms_per_step"])
for k in (j, j["train_b256"]):
    print([ (a[:22],round(b,4)) for a,b in k["stages_ms"]])
PY
