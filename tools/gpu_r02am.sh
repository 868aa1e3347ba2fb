set -x
mkdir -p gpurun_out
OLD=$PWD/slimg_b200/lib/libslimg_b200_old.so
( for c in c4a r3 r2 r15 c1 c3 thumbs; do
    AB=unroll2 timeout 120 python tools/time_config.py $c
    AB=old SLIMG_B200_LIB=$OLD timeout 120 python tools/time_config.py $c
  done ) > gpurun_out/configs_am.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_ratio_sweep.py tests/test_gpu_parity.py -m gpu -q > gpurun_out/pytest_gpu_am.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_am.log
