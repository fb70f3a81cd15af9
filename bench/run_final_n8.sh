#!/bin/bash
# Round-end multi-GPU validation on one 8-GPU box: pytest multi-GPU worker (2 and 8 ranks), then the full
# bench at N = 2, 4, 8 and the GroupBy phase split at 8.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -4
for N in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port $((29700+N)) bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.log 2>gpurun_out/bench_n$N.err
  echo "bench N=$N rc=$?"; tail -1 gpurun_out/bench_n$N.log | cut -c1-220
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 --master-port 29720 bench/groupby_phases.py 125000000 1048576 sparse > gpurun_out/groupby_phases_n8_sparse.log 2>&1; tail -1 gpurun_out/groupby_phases_n8_sparse.log | cut -c1-400
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 --master-port 29721 bench/terasort_phases.py > gpurun_out/phases_n8_final.log 2>&1; tail -1 gpurun_out/phases_n8_final.log | cut -c1-400
