#!/bin/bash
# usage: scripts/ncu_cnn.sh <tag> <batch>   -- ncu --set full of one LeNet train step's kernels (after the same command ran clean)
tag=$1; b=${2:-2048}
timeout 120 python scripts/run_kernel.py lenet_train $b 3 || exit 1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"conv_i2c|gemm_tf32|conv_strip" -s 16 -c 8 -f -o gpurun_out/${tag}_tc_b$b python scripts/run_kernel.py lenet_train $b 3 > gpurun_out/${tag}_tc_b$b.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"first_conv|head_|maxpool|adagrad|sgd|reduce_partials|act_bwd|split_params" -s 24 -c 12 -f -o gpurun_out/${tag}_simt_b$b python scripts/run_kernel.py lenet_train $b 3 > gpurun_out/${tag}_simt_b$b.log 2>&1
tail -n 2 gpurun_out/${tag}_tc_b$b.log; tail -n 2 gpurun_out/${tag}_simt_b$b.log
