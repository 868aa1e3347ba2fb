set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_paths_jpeg.py -m gpu -q > gpurun_out/pytest_gpu_k.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_k.log
( for c in thumb c3 c1 c4a r6 r55; do AB=default python tools/time_config.py $c; done
  AB=hfirst SLIMG_B200_RESIZE_PATH=hfirst python tools/time_config.py thumb
  AB=cols16 SLIMG_B200_STRIP_COLS=16 python tools/time_config.py thumb
) > gpurun_out/configs_k.txt 2>&1
