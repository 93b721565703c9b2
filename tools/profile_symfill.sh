#!/bin/bash
# `ncu --set full` capture of the shared-memory symbolic fill kernel on configs[1] (raw + source pages as CSV).
# The command is first run without ncu.  usage (under gpurun, repo root): bash tools/profile_symfill.sh
set -u
python tools/numeric_variants.py cfg2 "" > gpurun_out/plain_symfill.log 2>&1 || { echo "plain run failed"; tail -3 gpurun_out/plain_symfill.log; exit 1; }
ncu --set full --clock-control none --import-source on -k "regex:sym_fill_smem" -s 1 -c 1 -o gpurun_out/r02_symfill_v2 \
    python tools/numeric_variants.py cfg2 "" > gpurun_out/ncu_symfill.log 2>&1
tail -1 gpurun_out/ncu_symfill.log
ncu -i gpurun_out/r02_symfill_v2.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_sym_fill_smem_v2_raw.csv 2>/dev/null
ncu -i gpurun_out/r02_symfill_v2.ncu-rep --page source --csv --print-source cuda > gpurun_out/r02_ncu_source_sym_fill_smem_v2.csv 2>/dev/null
rm -f gpurun_out/r02_symfill_v2.ncu-rep
