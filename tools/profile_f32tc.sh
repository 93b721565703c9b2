#!/bin/bash
# `ncu --set full` capture of the float32 tensor-core kernel (k = 64 and k = 32 bench workloads), raw CSV only.
# Each command is first run without ncu.  usage (under gpurun, repo root): bash tools/profile_f32tc.sh
set -u
for k in 64 32; do
    python tools/f32_variants.py $k 3 > gpurun_out/plain_f32tcb$k.log 2>&1 || { echo "plain run k=$k failed"; tail -3 gpurun_out/plain_f32tcb$k.log; continue; }
    ncu --set full --clock-control none --import-source on -k "regex:f32_tcb" -s 1 -c 1 -o gpurun_out/r02_f32tcb$k \
        python tools/f32_variants.py $k 3 > gpurun_out/ncu_f32tcb$k.log 2>&1
    tail -1 gpurun_out/ncu_f32tcb$k.log
    ncu -i gpurun_out/r02_f32tcb$k.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_f32tcb${k}_raw.csv 2>/dev/null
    rm -f gpurun_out/r02_f32tcb$k.ncu-rep
done
