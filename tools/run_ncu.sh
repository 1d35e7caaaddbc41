#!/bin/bash
# ncu evidence for profiles/ (B200_PROFILING.md recipe).  Run on the GPU box AFTER the same bench command exited 0 without ncu.
#   tools/run_ncu.sh <tag>
TAG=$1
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity"
mkdir -p gpurun_out
$BENCH > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain bench failed"; exit 1; }
# (a) launch list: every kernel of the run, duration only
SPARKNET_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${TAG}_launches.csv $BENCH > gpurun_out/${TAG}_ncu_a.log 2>&1
# (b) the tensor group (contractions + their pack / reduce helpers) of two steady-state iterations, full set
TENSOR='regex:conv_window|wgrad_window|gemm_tc_kernel|flat_pack_kernel|s2d_pack_kernel|s2d_unpack|mat_reduce|wgrad_reduce|round_tf32|ip_round_bias|bias_grad_final|filter_pack_batch|splitk_reduce'
SPARKNET_GRAPH=0 ncu --set full --clock-control none -k "$TENSOR" -s 265 -c 72 -o gpurun_out/${TAG}_tensor $BENCH > gpurun_out/${TAG}_ncu_b.log 2>&1
ncu -i gpurun_out/${TAG}_tensor.ncu-rep --page raw --csv > gpurun_out/${TAG}_tensor_raw.csv 2>/dev/null
rm -f gpurun_out/${TAG}_tensor.ncu-rep      # the report is > 64 MiB: only the raw page travels back
# (c) the bandwidth-bound kernels of two iterations, full set
BW='regex:pool_|lrn_|sgd_update|widen_bytes|dropout_|relu_|softmax'
SPARKNET_GRAPH=0 ncu --set full --clock-control none -k "$BW" -s 100 -c 30 -o gpurun_out/${TAG}_bw $BENCH > gpurun_out/${TAG}_ncu_c.log 2>&1
ncu -i gpurun_out/${TAG}_bw.ncu-rep --page raw --csv > gpurun_out/${TAG}_bw_raw.csv 2>/dev/null
rm -f gpurun_out/${TAG}_bw.ncu-rep
ls -la gpurun_out/${TAG}_*
