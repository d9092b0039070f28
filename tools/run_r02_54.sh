set -x
mkdir -p gpurun_out
timeout 60 python tools/icp_time.py 16384 > gpurun_out/r02_54_plain_icp.log 2>&1 && timeout 200 ncu --set full --clock-control none --import-source on -k regex:icp_accumulate_kernel -s 3 -c 1 -f -o gpurun_out/r02_icp_accumulate python tools/icp_time.py 16384 > gpurun_out/ncu_icp54.log 2>&1
tail -2 gpurun_out/ncu_icp54.log; cat gpurun_out/r02_54_plain_icp.log | tail -1
