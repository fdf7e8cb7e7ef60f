#!/bin/bash
# One-GPU measurement set of a round: default bench line, reference arm, ncu launch list of the
# bench command and one `ncu --set full` capture of the two tensor-core kernels (each ncu pass only
# after the same command has run without it).  Everything lands in gpurun_out/.
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/r01_bench_1gpu.json 2> gpurun_out/bench_err.log
timeout 400 python bench.py --impl reference > gpurun_out/r01_bench_reference_arm.json 2>> gpurun_out/bench_err.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/r01_launches_bench_wide4096.csv python bench.py --steps 3 --warmup 3 --no-tiny --no-cpu-baseline \
  > gpurun_out/ncu_bench.log 2>&1
timeout 100 python tools/time_kernels.py > gpurun_out/time_kernels.log 2>&1
SPIN=lbwd timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mlp_fwd_tc_kernel|mlp_bwd_tc_kernel" \
  --launch-skip 60 --launch-count 3 -o gpurun_out/tc_kernels_v3 -f python tools/time_kernels.py > gpurun_out/ncu_tc.log 2>&1
tail -2 gpurun_out/ncu_tc.log
python tools/launch_summary.py gpurun_out/r01_launches_bench_wide4096.csv | tail -9
cat gpurun_out/time_kernels.log
cut -c1-300 gpurun_out/r01_bench_1gpu.json
