#!/bin/bash
# usage: scripts/scale_check.sh <tag> <ngpus>   -- data-parallel parity + bench at N GPUs; A/B: NCCL instead of the peer-memory exchange
tag=$1; n=$2
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port "$@"; }
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 tests/dp_check.py > gpurun_out/${tag}_dpcheck.log 2>&1; echo dpcheck rc=$?
grep dp_check gpurun_out/${tag}_dpcheck.log | sort -u
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 30 --warmup 5 > gpurun_out/${tag}_bench_n$n.json 2> gpurun_out/${tag}_bench_n$n.err; echo bench rc=$?
NC_DP_NO_P2P=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $n --steps 30 --warmup 5 --no-preprocess --no-cpu-baseline > gpurun_out/${tag}_bench_n${n}_nccl.json 2> gpurun_out/${tag}_bench_n${n}_nccl.err; echo nccl rc=$?
NC_DP_NO_OVERLAP=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus $n --steps 30 --warmup 5 --no-preprocess --no-cpu-baseline > gpurun_out/${tag}_bench_n${n}_nooverlap.json 2> gpurun_out/${tag}_bench_n${n}_nooverlap.err; echo nooverlap rc=$?
python - <<PY
import json
for f in ("${tag}_bench_n$n","${tag}_bench_n${n}_nccl","${tag}_bench_n${n}_nooverlap"):
    try:
        j=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
        print(f, j["ms_per_step"], j["value"], j["config"].get("parallelism")[:40], {k:(v.get("ms_per_step") if isinstance(v,dict) else v) for k,v in j.items() if k.startswith("train_")}, (j.get("dp_parity") or {}).get("error_relative_to_update"))
    except Exception as e: print(f, "ERR", e)
PY
