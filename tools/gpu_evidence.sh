#!/bin/bash
# One pass of the round's GPU evidence (run on the GPU box through gpurun; writes under gpurun_out/evidence/):
#   l2_*.csv          ssn_score_kernel<false> over every column of the configs[1] batch with the persisting L2 window on coef+lhet
#                     on and off (SSN_NO_L2_WINDOW=1): duration, L2 hit rate, L2 read sectors from the SMs, DRAM bytes
#   ncu_*.ncu-rep     full-set captures of the pre-filter kernel, the full model, and the deep (panel) path
#   sanitizer_*.log   compute-sanitizer memcheck / racecheck / synccheck over tools/sanitize_gpu.py
set -u
OUT=gpurun_out/evidence
mkdir -p $OUT
M="gpu__time_duration.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum,lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum,lts__t_sectors_op_read.sum,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__t_sector_hit_rate.pct,smsp__inst_executed.sum"
ncu --metrics $M --clock-control none -k regex:ssn_score_kernel -s 3 -c 3 --csv --log-file $OUT/l2_window_on.csv python tools/kexp.py full > $OUT/l2_window_on.out 2>&1
SSN_NO_L2_WINDOW=1 ncu --metrics $M --clock-control none -k regex:ssn_score_kernel -s 3 -c 3 --csv --log-file $OUT/l2_window_off.csv python tools/kexp.py full > $OUT/l2_window_off.out 2>&1
python tools/kexp.py full > $OUT/time_window_on.json 2>/dev/null
SSN_NO_L2_WINDOW=1 python tools/kexp.py full > $OUT/time_window_off.json 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:ssn_prefilter_kernel -s 5 -c 1 -o $OUT/ncu_prefilter -f python tools/kexp.py wgs > $OUT/ncu_prefilter.out 2>&1
ncu --set full --clock-control none --import-source on -k regex:ssn_score_kernel -s 3 -c 1 -o $OUT/ncu_full_model -f python tools/kexp.py full > $OUT/ncu_full_model.out 2>&1
KEXP_REPS=2 ncu --set full --clock-control none --import-source on -k regex:ssn_score_kernel -s 3 -c 1 -o $OUT/ncu_panel -f python tools/kexp.py panel > $OUT/ncu_panel.out 2>&1
for tool in memcheck racecheck synccheck; do
    timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_gpu.py > $OUT/sanitizer_$tool.log 2>&1
    echo "exit code $?" >> $OUT/sanitizer_$tool.log
    tail -3 $OUT/sanitizer_$tool.log
done
ls -la $OUT
