#!/bin/bash
# Round-2 ncu evidence (run on the GPU box: `gpurun -- bash profiles/capture_r2.sh`).  Every capture runs only after the same
# command exited 0 without ncu; nothing printed under ncu is a bench value.  Outputs land in gpurun_out/ and are
# summarised into profiles/ by profiles/summarize.py / profiles/make_kernel_profile.py (see profiles/README.md).
set -x
O=gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-side-metrics --strong-spp 0"
$B > $O/r2_plain.json 2> $O/r2_plain.err || exit 1
# 1. launch list of the bench command (kernel shares of the step)
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r2_launches.csv $B > $O/r2_launches.log 2>&1
# 2. per-launch counters of the three hot kernels over one bench run (traffic, instructions, lanes)
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active \
    --clock-control none -k regex:"k_trace|k_shade|k_shadow|k_classify" -c 200 --csv --log-file $O/r2_counters.csv $B > $O/r2_counters.log 2>&1
# 3. full sets with source: rgb step (closest-hit primary + 2nd bounce, shade, shadow), spectral shade, gradient shade
python tools/prof_run.py rgb_time > /dev/null 2>&1 && ncu --set full --import-source on --clock-control none -k regex:"k_trace|k_shade|k_shadow" -c 6 -o $O/r2_rgb python tools/prof_run.py rgb_time > $O/r2_rgb.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_shade -c 1 -o $O/r2_spectral python tools/prof_run.py spectral > $O/r2_spectral.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_shade_grad -c 1 -o $O/r2_grad python tools/prof_run.py grad > $O/r2_grad.log 2>&1
# text summaries are made here (the .ncu-rep files together exceed what gpurun copies back); only the rgb report travels
for r in r2_rgb r2_spectral r2_grad; do
  python profiles/summarize.py raw $O/$r.ncu-rep > $O/${r}_summary.txt 2>&1
  python profiles/summarize.py hot $O/$r.ncu-rep 0 40 > $O/${r}_hot.txt 2>&1
done
python profiles/summarize.py hot $O/r2_rgb.ncu-rep 3 40 > $O/r2_rgb_hot_bounce2.txt 2>&1
python profiles/summarize.py hot $O/r2_rgb.ncu-rep 1 40 > $O/r2_rgb_hot_shade.txt 2>&1
python profiles/summarize.py bysource $O/r2_rgb.ncu-rep "k_shade" 30 > $O/r2_rgb_shade_by_source.txt 2>&1
python profiles/summarize.py bysource $O/r2_rgb.ncu-rep "k_trace" 30 > $O/r2_rgb_trace_by_source.txt 2>&1
python profiles/summarize.py launches $O/r2_launches.csv > $O/r2_launches_summary.txt 2>&1
python profiles/make_kernel_profile.py $O/r2_counters.csv $O/r2_plain.json $O/r2_kernel_profile.json > $O/r2_kernel_profile.log 2>&1
rm -f $O/r2_spectral.ncu-rep $O/r2_grad.ncu-rep
ls -la $O
