#!/bin/bash
# usage: scripts/p2p_n2.sh  -- 2 GPUs: A/B of the gradient exchange (CTA count of the peer-memory kernel, NCCL, no overlap)
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 30 --warmup 5 --no-preprocess --no-cpu-baseline > gpurun_out/p2p_$tag.json 2> gpurun_out/p2p_$tag.err; echo $tag rc=$?; }
run default X=1
run ctas16 NC_P2P_CTAS=16
run ctas96 NC_P2P_CTAS=96
run nooverlap NC_DP_NO_OVERLAP=1
run nccl NC_DP_NO_P2P=1
python - <<'PY'
import json
for f in ("default","ctas16","ctas96","nooverlap","nccl"):
    try:
        j=json.loads(open("gpurun_out/p2p_%s.json"%f).read().strip().splitlines()[-1])
        print(f, round(j["ms_per_step"],4), {k:round(v.get("ms_per_step"),4) for k,v in j.items() if k.startswith("train_")})
    except Exception as e: print(f, "ERR", e)
PY
