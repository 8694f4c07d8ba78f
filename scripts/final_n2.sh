mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dp_check.py > gpurun_out/r2f_dpcheck_n2.log 2>&1; echo dpcheck rc=$?
grep dp_check gpurun_out/r2f_dpcheck_n2.log | sort -u
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/r2f_bench_n2.json 2> gpurun_out/r2f_bench_n2.err; echo bench rc=$?
head -c 1500 gpurun_out/r2f_bench_n2.json
