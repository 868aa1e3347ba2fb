set -x
mkdir -p gpurun_out
OLD=$PWD/slimg_b200/lib/libslimg_b200_old.so
( for rep in 1 2; do for c in c4a r3 r2 r15 thumbs c3 c1; do
    AB=new timeout 120 python tools/time_config.py $c
    AB=old SLIMG_B200_LIB=$OLD timeout 120 python tools/time_config.py $c
  done; done ) > gpurun_out/configs_ap.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_ap.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_ap.log
