# Run on the GPU box (gpurun -- 'bash profiles/capture.sh'): the raw captures that
# profiles/make_summaries.py turns into the tracked summaries.  Numbers printed under ncu are
# never bench values.
TAG=${1:-r02}
B="python bench.py --gpus 1 --steps 2 --warmup 1 --no-cpu-baseline --no-extra --e2e-steps 1 --scan-steps 2"
$B > gpurun_out/${TAG}_prof_bench.json 2> gpurun_out/${TAG}_prof_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:^(kinetics_|stencil_|correct_kernel|permeate_|scatter_|degrade_kernel|pixel_index|early_exit_verdict)' -c 200 \
    --csv --log-file gpurun_out/${TAG}_launch_list_raw.csv $B > gpurun_out/${TAG}_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:kinetics_lanes -s 2 -c 1 -o gpurun_out/${TAG}_kin_cached -f $B > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:kinetics_lanes -s 2 -c 1 -o gpurun_out/${TAG}_kin_scan -f $B --no-list-cache > /dev/null 2>&1
ls -la gpurun_out
