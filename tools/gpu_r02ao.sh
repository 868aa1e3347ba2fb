set -x
mkdir -p gpurun_out
OLD=$PWD/slimg_b200/lib/libslimg_b200_old.so
( for rep in 1 2; do for c in c4b up2; do
    AB=new timeout 120 python tools/time_config.py $c
    AB=old SLIMG_B200_LIB=$OLD timeout 120 python tools/time_config.py $c
  done; done
  for c in r15 r2; do
    AB=tile_new SLIMG_B200_RESIZE_PATH=tile timeout 120 python tools/time_config.py $c
    AB=tile_old SLIMG_B200_RESIZE_PATH=tile SLIMG_B200_LIB=$OLD timeout 120 python tools/time_config.py $c
  done ) > gpurun_out/configs_ao.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_ao.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_ao.log
