set -x
mkdir -p gpurun_out
( for d in 11 27 0 16 24 8; do AB=tc_debug$d SLIMG_B200_VF_DEBUG=$d SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3; done
  for d in 11 27; do AB=tc_debug${d}_c4a SLIMG_B200_VF_DEBUG=$d SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c4a; done
  for d in 3 7; do AB=vf_debug$d SLIMG_B200_VF_DEBUG=$d python tools/time_config.py c3; done
  SLIMG_B200_VF_DEBUG=16 SLIMG_B200_RESIZE_PATH=tc python tools/check_one.py 6000 4000 800; SLIMG_B200_RESIZE_PATH=tc python tools/check_one.py 6000 4000 800
) > gpurun_out/configs_d.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_paths_jpeg.py tests/test_gpu_uniffi.py -m gpu -q -x > gpurun_out/pytest_gpu_d.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_d.log
