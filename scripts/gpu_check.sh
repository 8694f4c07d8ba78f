#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench (both arms), then -- only if everything passed -- the ncu launch list.
# usage: scripts/gpu_check.sh <tag> [ncu_kernel_regex]
tag=${1:-run}
kre=${2:-}
mkdir -p gpurun_out
nvidia-smi > gpurun_out/${tag}_smi.txt 2>&1
(nproc; free -g; lscpu | head -20) > gpurun_out/${tag}_host.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/${tag}_pytest.log 2>&1
rc=$?
tail -8 gpurun_out/${tag}_pytest.log
if [ $rc -ne 0 ]; then echo "pytest failed rc=$rc"; fi
timeout 300 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1 || { tail -20 gpurun_out/${tag}_smoke.log; }
timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || { tail -20 gpurun_out/${tag}_bench.err; }
head -c 3000 gpurun_out/${tag}_bench.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err
head -c 600 gpurun_out/${tag}_bench_ref.json
# ncu: launch list of a short train-only bench
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-preprocess --no-cpu-baseline --no-graph > gpurun_out/${tag}_ncu_list.log 2>&1
if [ -n "$kre" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$kre -s 6 -c 3 -f -o gpurun_out/${tag}_prof \
    python bench.py --steps 3 --warmup 3 --no-preprocess --no-cpu-baseline --no-graph > gpurun_out/${tag}_ncu_full.log 2>&1
fi
echo done
