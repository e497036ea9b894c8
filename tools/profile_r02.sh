#!/bin/bash
# Round-2 ncu evidence (run on the GPU box through gpurun; every command below has exited 0 without ncu before).  Outputs go to gpurun_out/.
#   1. launch list of the bench command (device time per launch, cold caches, serialised)
#   2. --set full of six consecutive wave launches of the `waves` block: once with ncu's default cache control (cold L2 per launch) and
#      once with --cache-control none (what the timed run sees: the scene stays in L2)
#   3. --set full of the frame's other kernels (shade, accumulate, generate) and of the scene-update kernels (refit walk, top-level build)
set -x
B="python bench.py --steps 2 --warmup 1 --wave-steps 2 --no-cpu-baseline --no-e2e --no-check"
$B > gpurun_out/r02_plain.json 2> gpurun_out/r02_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_launches.csv $B > gpurun_out/r02_ncu_launches.log 2>&1
# waves block: 3 warm-up passes x 6 waves come first
ncu --set full --clock-control none --import-source on -k regex:traceKernel -s 18 -c 6 -f -o gpurun_out/r02_trace $B --no-refit > gpurun_out/r02_ncu_trace.log 2>&1
ncu --set full --clock-control none --cache-control none --import-source on -k regex:traceKernel -s 18 -c 6 -f -o gpurun_out/r02_trace_warm $B --no-refit > gpurun_out/r02_ncu_trace_warm.log 2>&1
# the frame's own kernels: first pass of the first frame after the waves block
ncu --set full --clock-control none --import-source on -k regex:'shadeKernel|accumulateKernel|generateEyeRays' -c 8 -f -o gpurun_out/r02_frame $B --no-waves --no-refit > gpurun_out/r02_ncu_frame.log 2>&1
# scene update: C4 (1000 instances x 100K-triangle mesh: refit + top-level build per frame) and the 1M-triangle refit
python tools/bench_configs.py refit c4 --frames 3 > gpurun_out/r02_cfg_plain.jsonl 2> gpurun_out/r02_cfg_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:'refitWalkKernel|tlasBuildKernel|refitParents' -c 10 -f -o gpurun_out/r02_update python tools/bench_configs.py refit c4 --frames 3 > gpurun_out/r02_ncu_update.log 2>&1
ls -la gpurun_out/*.ncu-rep
# the default mesh builder (binned SAH on the device) and the refit as it is at the end of the round
python tools/probes/build_once.py > gpurun_out/r02_build_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'top.*Kernel|sahb.*Kernel|binaryBoxesKernel|collapse.*Kernel|refitWalkKernel|primBoundsKernel' -c 120 -f -o gpurun_out/r02_build python tools/probes/build_once.py > gpurun_out/r02_ncu_build.log 2>&1
python tools/ncu_kernel_summary.py gpurun_out/r02_build.ncu-rep --json gpurun_out/r02_ncu_build.json > gpurun_out/r02_ncu_build.txt
