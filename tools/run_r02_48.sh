mkdir -p gpurun_out
cp 3dbodyanimation_b200/lib3dba_b200.so /tmp/lib_keep.so
for v in gpurun_var/lib_*.so; do
  cp $v 3dbodyanimation_b200/lib3dba_b200.so
  echo "== $v"
  for i in 1 2; do timeout 60 python tools/prof.py forward 65536 f32 2>&1 | tail -1; done; timeout 120 python -m pytest tests/test_forward_gpu.py -x -q -m gpu 2>&1 | tail -1
done
cp /tmp/lib_keep.so 3dbodyanimation_b200/lib3dba_b200.so
