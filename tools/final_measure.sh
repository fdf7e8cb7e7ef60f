#!/bin/bash
# One-GPU measurement set of a round (round tag = $1, default r02): default bench line, reference arm, then -- each ncu
# pass only after the same command has run without it -- the ncu launch list of the bench command and one
# `ncu --set full` capture of the step's kernels.  `fh` as $2 does the same for the wide-input workload instead.
# Everything lands in gpurun_out/; tools/ncu_summary.py, tools/launch_summary.py and tools/make_traffic_json.py turn
# the reports into the text files under profiles/.
R=${1:-r02}
mkdir -p gpurun_out
if [ "$2" = "fh" ]; then
  timeout 400 python bench.py --workload wideFH4096 > gpurun_out/${R}_bench_wideFH4096_1gpu.json 2> gpurun_out/bench_fh_err.log
  timeout 300 python bench.py --workload wide8192 > gpurun_out/${R}_bench_wide8192_1gpu.json 2>> gpurun_out/bench_fh_err.log
  B200DS_BENCH_SETTLE=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"fh_" \
    --launch-skip 40 --launch-count 12 -o gpurun_out/${R}_fh_kernels -f python bench.py --workload wideFH4096 --steps 2 --warmup 3 --no-tiny --no-cpu-baseline \
    > gpurun_out/ncu_fh.log 2>&1
  tail -2 gpurun_out/ncu_fh.log
  cut -c1-300 gpurun_out/${R}_bench_wideFH4096_1gpu.json
  exit 0
fi
timeout 600 python bench.py > gpurun_out/${R}_bench_wide4096_1gpu.json 2> gpurun_out/bench_err.log
timeout 400 python bench.py --impl reference > gpurun_out/${R}_bench_reference_arm.json 2>> gpurun_out/bench_err.log
timeout 100 python tools/time_kernels.py > gpurun_out/time_kernels.log 2>&1
B200DS_BENCH_SETTLE=5 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/${R}_launches_bench_wide4096.csv python bench.py --steps 3 --warmup 3 --no-tiny --no-cpu-baseline \
  > gpurun_out/ncu_bench.log 2>&1
B200DS_BENCH_SETTLE=5 timeout 600 ncu --set full --clock-control none --import-source on \
  -k regex:"gen_batch_kernel|mlp_fwd_tc_kernel|mlp_bwd_tc_kernel|mlp_update_tc_kernel" --launch-skip 24 --launch-count 8 \
  -o gpurun_out/${R}_tc_kernels -f python bench.py --steps 3 --warmup 3 --no-tiny --no-cpu-baseline > gpurun_out/ncu_tc.log 2>&1
tail -2 gpurun_out/ncu_tc.log
python tools/launch_summary.py gpurun_out/${R}_launches_bench_wide4096.csv | tail -12
cat gpurun_out/time_kernels.log
cut -c1-300 gpurun_out/${R}_bench_wide4096_1gpu.json
