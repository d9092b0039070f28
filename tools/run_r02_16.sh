set -x
mkdir -p gpurun_out
python tools/prof.py forward 65536 f32 > gpurun_out/r02_plain_fwd.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:smpl_lbs_tc_kernel -s 3 -c 1 -o gpurun_out/r02_lbs_tc_final python tools/prof.py forward 65536 f32 > gpurun_out/ncu_lbs.log 2>&1
tail -2 gpurun_out/ncu_lbs.log; cat gpurun_out/r02_plain_fwd.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_n1_b.json 2> gpurun_out/r02_bench_n1_b.err; echo "bench rc $?"; tail -c 300 gpurun_out/r02_bench_n1_b.err
