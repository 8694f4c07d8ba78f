#!/bin/bash
# usage: scripts/scale_check.sh <tag> <ngpus>   -- data-parallel parity + bench at N GPUs, with A/B switches
tag=$1; n=$2
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 tests/dp_check.py > gpurun_out/${tag}_dpcheck.log 2>&1; echo dpcheck rc=$?
grep dp_check gpurun_out/${tag}_dpcheck.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 30 --warmup 5 > gpurun_out/${tag}_bench_n$n.json 2> gpurun_out/${tag}_bench_n$n.err; echo bench rc=$?
grep "^\[bench" gpurun_out/${tag}_bench_n$n.err | tail -3
NC_DP_NO_EARLY_APPLY=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $n --steps 30 --warmup 5 --no-preprocess > gpurun_out/${tag}_bench_n${n}_noearly.json 2> gpurun_out/${tag}_bench_n${n}_noearly.err; echo noearly rc=$?
NCCL_MAX_CTAS=8 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $n --steps 30 --warmup 5 --no-preprocess > gpurun_out/${tag}_bench_n${n}_cta8.json 2> gpurun_out/${tag}_bench_n${n}_cta8.err; echo cta8 rc=$?
NC_DP_NO_OVERLAP=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus $n --steps 30 --warmup 5 --no-preprocess > gpurun_out/${tag}_bench_n${n}_nooverlap.json 2> gpurun_out/${tag}_bench_n${n}_nooverlap.err; echo nooverlap rc=$?
