set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 8 --warmup 5 --no-cpu-baseline --no-e2e --warm-seconds 0 --no-clocks --configs 1"
$CMD > gpurun_out/plain1.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:classify_kernel -s 6 -c 3 -o gpurun_out/r02_prof_cfg1 $CMD > gpurun_out/ncu2.log 2>&1
python tools/prof_big.py 3 > gpurun_out/r02_big_plain.log 2>&1 && ncu --set full --clock-control none --cache-control none -k regex:classify_kernel -s 1 -c 2 -o gpurun_out/r02_prof_big python tools/prof_big.py 3 > gpurun_out/ncu3.log 2>&1
M="dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum,gpu__time_duration.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,sm__inst_executed.avg.per_cycle_elapsed,launch__registers_per_thread,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"
for c in cfg3 cfg4 cfg5; do
  SFC_FLAGS=2 SFC_REPS=2 python tools/prof_cfg.py $c 6 > gpurun_out/plain_$c.log 2>&1 && SFC_FLAGS=2 SFC_REPS=2 ncu --metrics $M --clock-control none --cache-control none -k regex:classify_kernel -s 8 -c 3 --csv --log-file gpurun_out/r02_ncu_$c.csv python tools/prof_cfg.py $c 6 > gpurun_out/ncu_$c.log 2>&1
done
ls -la gpurun_out | tail -20
