mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dp_check.py > gpurun_out/p2p_dpcheck_n2.log 2>&1; echo dpcheck rc=$?
grep dp_check gpurun_out/p2p_dpcheck_n2.log; tail -5 gpurun_out/p2p_dpcheck_n2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 30 --warmup 5 --no-preprocess > gpurun_out/p2p_bench_n2.json 2> gpurun_out/p2p_bench_n2.err; echo bench rc=$?
NC_DP_NO_P2P=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 30 --warmup 5 --no-preprocess > gpurun_out/p2p_bench_n2_nccl.json 2> gpurun_out/p2p_bench_n2_nccl.err; echo nccl rc=$?
python - <<'PY'
import json
for f in ("p2p_bench_n2","p2p_bench_n2_nccl"):
    try:
        j=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
        print(f, j["ms_per_step"], j["value"], j["config"].get("parallelism"), {k:(v.get("ms_per_step") if isinstance(v,dict) else v) for k,v in j.items() if k.startswith("train_")}, j.get("dp_parity"))
    except Exception as e: print(f, "ERR", e)
PY
