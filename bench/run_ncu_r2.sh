#!/bin/bash
# ncu evidence for the final kernels (one GPU).  The program runs once WITHOUT ncu first; the report is
# exported to CSV on the box (the .ncu-rep is larger than what gpurun copies back).
mkdir -p gpurun_out
timeout 300 python bench/profile_r2.py > gpurun_out/profile_r2.log 2>&1 || { echo "target failed"; tail -5 gpurun_out/profile_r2.log; exit 1; }
tail -1 gpurun_out/profile_r2.log
KERN='regex:sh_scatter_tma|sh_count|rs_onesweep|gather_rows_fixed|agg_update_kernel|km_assign_pair|km_certify|tok_write|dryad_fused|fix_ties32_keys|sh_scatter'
timeout 900 ncu --set full --clock-control none --import-source on -k "$KERN" -c 40 -f -o /tmp/ncu_r2 python bench/profile_r2.py > gpurun_out/ncu_r2.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_r2.log
ncu -i /tmp/ncu_r2.ncu-rep --page raw --csv > gpurun_out/ncu_r2_raw.csv 2>/dev/null
ncu -i /tmp/ncu_r2.ncu-rep --page details --csv > gpurun_out/ncu_r2_details.csv 2>/dev/null
ncu -i /tmp/ncu_r2.ncu-rep --page source --csv -k regex:km_assign_pair -c 1 > gpurun_out/ncu_r2_source_km_assign_pair.csv 2>/dev/null
ncu -i /tmp/ncu_r2.ncu-rep --page source --csv -k regex:sh_scatter_tma -c 1 > gpurun_out/ncu_r2_source_sh_scatter_tma.csv 2>/dev/null
ls -la gpurun_out/
