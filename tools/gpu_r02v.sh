set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ratio_sweep.py -m gpu -q > gpurun_out/pytest_sweep.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_sweep.log
( for c in r2 r15 r3; do for d in 0 1 2 3; do AB=vf_debug$d SLIMG_B200_VF_DEBUG=$d timeout 120 python tools/time_config.py $c; done; done
  for c in r2 r15; do for sc in 32 48; do AB=cols$sc SLIMG_B200_STRIP_COLS=$sc timeout 120 python tools/time_config.py $c; done; done
  for c in r2 r15; do AB=tile SLIMG_B200_RESIZE_PATH=tile timeout 120 python tools/time_config.py $c; done
) > gpurun_out/configs_v.txt 2>&1
