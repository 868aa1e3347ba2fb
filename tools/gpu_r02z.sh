set -x
mkdir -p gpurun_out
L=$PWD/slimg_b200/lib
( for c in c3 c3conv c4a c1 r6 r55 r3 r2 r15 thumb thumbs; do
    AB=new timeout 120 python tools/time_config.py $c
  done
  for c in c3 c3conv; do AB=old SLIMG_B200_LIB=$L/libslimg_b200_old.so timeout 120 python tools/time_config.py $c; done
) > gpurun_out/configs_z.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_z.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_z.log
