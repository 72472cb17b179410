# One gpurun call: the bench command plain, then its ncu launch list and one --set full capture of the dominant kernel
# (B200_PROFILING.md recipe).  Outputs under gpurun_out/.
set -x
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-reference-layout"
K='regex:dense_|flag_|scan_|reduce_|sortkey_|copy_words|partition_|kv_|hash_|for_each'
$CMD > gpurun_out/r02_profile_plain.json 2> gpurun_out/profile_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 300 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:dense_count_kernel -s 3 -c 1 -o gpurun_out/r02_emit_count $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
