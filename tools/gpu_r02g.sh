set -x
mkdir -p gpurun_out
( for v in 5 6 8 9 0; do SLIMG_B200_JPEG_VAR=$v python tools/time_jpeg.py; done ) > gpurun_out/time_jpeg_g.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_paths_jpeg.py -m gpu -q -x > gpurun_out/pytest_gpu_g.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_g.log
( for c in c3 c1 c4a r3; do AB=tc SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py $c; done
  for d in 32 33 43; do AB=tc_debug$d SLIMG_B200_VF_DEBUG=$d SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3; done
  for vb in 2 4; do AB=tc_vb$vb SLIMG_B200_TC_VBUFS=$vb SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3; done
  AB=tc_pf0 SLIMG_B200_TC_PREFETCH=0 SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3
  AB=tc_3d SLIMG_B200_VF_DEBUG=16 SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3
) > gpurun_out/configs_g.txt 2>&1
