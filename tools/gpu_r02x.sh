set -x
mkdir -p gpurun_out
OLD=$PWD/slimg_b200/lib/libslimg_b200_old.so
( for c in c3 c3conv c4a c1 r6 r3 r2 r15 thumb thumbs; do
    AB=new timeout 120 python tools/time_config.py $c
    AB=old SLIMG_B200_LIB=$OLD timeout 120 python tools/time_config.py $c
  done
  for c in c3 c1 r6 r55 thumb; do for pf in 0 8; do AB=new_pf$pf SLIMG_B200_PREFETCH_STEPS=$pf timeout 120 python tools/time_config.py $c; done; done
) > gpurun_out/configs_x.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_x.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_x.log
