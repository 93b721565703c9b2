#!/bin/bash
# One `ncu --set full` capture per kernel family, each on the workload where that kernel matters
# (run under gpurun from the repo root; reports land in gpurun_out/r02_*.ncu-rep).  Every command is
# first run without ncu (guides/B200_PROFILING.md).  usage: bash tools/profile_all.sh
set -u
run() { # name, kernel regex, skip, count, command...
    local name=$1 regex=$2 skip=$3 count=$4; shift 4
    "$@" > gpurun_out/plain_$name.log 2>&1 || { echo "plain run of $name failed"; tail -3 gpurun_out/plain_$name.log; return; }
    ncu --set full --clock-control none --import-source on -k "regex:$regex" -s $skip -c $count -o gpurun_out/r02_$name "$@" > gpurun_out/ncu_$name.log 2>&1
    tail -1 gpurun_out/ncu_$name.log
    # only the exports travel back (the reports together exceed what gpurun copies)
    ncu -i gpurun_out/r02_$name.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_${name}_raw.csv 2>/dev/null
    rm -f gpurun_out/r02_$name.ncu-rep
}
run bigtile16 numeric_bigtile 2 1 python tools/numeric_variants.py k16 ""
run bigtile32 numeric_bigtile 2 1 python tools/numeric_variants.py k32 ""
run rows4_dense "numeric_rows|sym_fill_dense" 4 2 python tools/numeric_variants.py chaintop ""
run f32simt f32_simt 1 1 python tools/f32_variants.py 64 3
run f32tc32 f32_tc 1 1 python tools/f32_variants.py 32 3
run rmat_sym "sym_fill_tiny|sym_fill_global|hv_|sym_count_heavy" 0 12 python tools/rmat_trace.py
run merge "tile_nonzero|compact_tiles|checksum_kernel|add_rowptr|remap_rowptr" 0 8 python tools/merge_phase_bench.py
