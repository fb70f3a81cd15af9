#!/bin/bash
# ncu evidence for the final kernels (one GPU).  The program runs once WITHOUT ncu first.
mkdir -p gpurun_out
timeout 300 python bench/profile_r2.py > gpurun_out/profile_r2.log 2>&1 || { echo "target failed"; tail -5 gpurun_out/profile_r2.log; exit 1; }
tail -1 gpurun_out/profile_r2.log
KERN='regex:sh_scatter_tma|sh_count|rs_onesweep|gather_rows_fixed|agg_update_kernel|km_assign_pair|km_certify|tok_write|dryad_fused|fix_ties32_keys|sh_scatter'
timeout 900 ncu --set full --clock-control none --import-source on -k "$KERN" -c 40 -f -o gpurun_out/ncu_r2 python bench/profile_r2.py > gpurun_out/ncu_r2.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_r2.log
ls -la gpurun_out/ncu_r2.ncu-rep
