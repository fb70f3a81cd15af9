#!/bin/bash
# Multi-GPU development run: correctness worker, TeraSort phase splits for a few settings, kernel-path bench.
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
DRYAD_SHUFFLE_VERIFY=1 timeout 600 $TR --master-port 29611 tests/mgpu/spmd_worker.py --mode gpu > gpurun_out/mgpu$N.log 2>&1; echo "mgpu rc=$?"
tail -1 gpurun_out/mgpu$N.log | cut -c1-300
for cfg in ${CFGS:-"148:2" "148:4" "296:2" "74:2" "148:1" "0:2"}; do
  c=${cfg%%:*}; p=${cfg##*:}
  DRYAD_SORT_PAYLOAD_CTAS=$c DRYAD_SORT_PASSES=$p timeout 300 $TR --master-port 29612 bench/terasort_phases.py > gpurun_out/phases_n${N}_c${c}_p$p.log 2>&1; echo "ctas=$c passes=$p rc=$?"; tail -1 gpurun_out/phases_n${N}_c${c}_p$p.log | cut -c1-400
done
DRYAD_SORT_SPLIT=0 timeout 300 $TR --master-port 29613 bench/terasort_phases.py > gpurun_out/phases_n${N}_nosplit.log 2>&1; echo "nosplit rc=$?"; tail -1 gpurun_out/phases_n${N}_nosplit.log | cut -c1-400
timeout 300 $TR --master-port 29614 bench.py --gpus $N --steps 10 --warmup 3 --skip-e2e --skip-groupby > gpurun_out/bench_n$N.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_n$N.log | cut -c1-300
