set -x
mkdir -p gpurun_out
( for v in 5 6 7 8 9 0; do SLIMG_B200_JPEG_VAR=$v python tools/time_jpeg.py; done ) > gpurun_out/time_jpeg_f.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_paths_jpeg.py -m gpu -q -x -k "jpeg or Jpeg or JPEG or config or constant" > gpurun_out/pytest_gpu_f.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_f.log
( for d in 32 43 33 34 40; do AB=tc_debug$d SLIMG_B200_VF_DEBUG=$d SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3; done
  for c in c4b up2 thumbs; do AB=default python tools/time_config.py $c; done
  AB=tile64x64 SLIMG_B200_TILE_W=64 python tools/time_config.py c4b
) > gpurun_out/configs_f.txt 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jpeg_frontend -s 1 -c 1 -o gpurun_out/k3_tensor_full python tools/time_jpeg.py > gpurun_out/ncu_k3.log 2>&1
