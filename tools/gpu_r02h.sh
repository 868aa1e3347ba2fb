set -x
mkdir -p gpurun_out
SLIMG_B200_RESIZE_PATH=tc timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "resize or config or chain or convert or guard or batch" > gpurun_out/pytest_gpu_h.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_h.log
( for c in c3 c1 c4a r3; do AB=tc SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py $c; done
  for d in 32 33 43; do AB=tc_debug$d SLIMG_B200_VF_DEBUG=$d SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3; done
  AB=tc_slots4 SLIMG_B200_TC_SLOTS=4 SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3
  SLIMG_B200_JPEG_VAR=8 python tools/time_jpeg.py
) > gpurun_out/configs_h.txt 2>&1
