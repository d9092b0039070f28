set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_forward_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02_49_pytest_forward.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r02_49_pytest_forward.log
for i in 1 2; do timeout 120 python tools/prof.py forward 65536 f32; done 2>&1 | tail -2
python tools/prof.py forward 65536 f32 > gpurun_out/r02_49_plain_fwd.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:smpl_lbs_tc_kernel -s 3 -c 1 -f -o gpurun_out/r02_lbs_tc_v2 python tools/prof.py forward 65536 f32 > gpurun_out/ncu_lbs49.log 2>&1
tail -2 gpurun_out/ncu_lbs49.log
