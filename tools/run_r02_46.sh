set -x
mkdir -p gpurun_out
python tools/prof.py forward 65536 f32 > gpurun_out/r02_46_plain_fwd.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:smpl_lbs_tc_kernel -s 3 -c 1 -f -o gpurun_out/r02_lbs_tc_ffma2 python tools/prof.py forward 65536 f32 > gpurun_out/ncu_lbs46.log 2>&1
tail -2 gpurun_out/ncu_lbs46.log; cat gpurun_out/r02_46_plain_fwd.log
