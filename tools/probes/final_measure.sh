#!/bin/bash
# Probe (not product): the measurement pass behind profiles/r02_*: bench lines first (no profiler), then the same frame under ncu.
#   gpurun --timeout 1500 -- 'bash tools/probes/final_measure.sh'
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python bench.py > $O/r02_bench.json 2> $O/r02_bench.err
timeout 300 python bench.py --impl reference > $O/r02_bench_reference.json 2> $O/r02_bench_reference.err
timeout 300 python tools/probes/pipeline_probe.py 40 $O/r02_pipeline_timeline.txt > $O/r02_pipeline_probe.txt 2>&1
LSB_HOST_MODE=1 timeout 300 python tools/probes/pipeline_probe.py 40 - > $O/r02_pipeline_probe_hostmode1.txt 2>&1
timeout 300 python tools/probes/profile_frame.py > $O/r02_profile_frame.txt 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/r02_launches.csv \
  python tools/probes/profile_frame.py > $O/r02_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -o $O/r02_frame -f \
  python tools/probes/profile_frame.py > $O/r02_ncu_full.log 2>&1
ncu -i $O/r02_frame.ncu-rep --page raw --csv > $O/r02_frame_raw.csv 2> /dev/null
rm -f $O/r02_frame.ncu-rep      # ~100 MB: gpurun brings back at most 64 MiB, the raw page is what the summaries are made from
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks.mem,clocks_throttle_reasons.active,power.draw --format=csv > $O/r02_gpu.txt
tail -c 300 $O/r02_bench.json; echo; tail -n 3 $O/r02_ncu_full.log; ls -la $O; du -sh $O
