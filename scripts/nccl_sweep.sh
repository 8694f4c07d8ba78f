#!/bin/bash
n=${1:-8}
mkdir -p gpurun_out
probe() { env "$@" timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29520 scripts/nccl_probe.py 2>/dev/null | grep nccl_probe; }
probe X=1
probe NCCL_ALGO=NVLS
probe NCCL_ALGO=Ring
probe NCCL_ALGO=Tree
probe NCCL_PROTO=Simple
probe NCCL_PROTO=LL128
probe NCCL_MAX_NCHANNELS=8
probe NCCL_MIN_NCHANNELS=32
probe NCCL_NVLS_ENABLE=0
NCCL_DEBUG=INFO timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29520 scripts/nccl_probe.py 2>&1 | grep -i -E "nvls|channels|algo" | head -20
