set -x
mkdir -p gpurun_out
./tools/tma_probe 880 > gpurun_out/tma_probe.txt 2>&1
./tools/tma_probe 656 >> gpurun_out/tma_probe.txt 2>&1
( for c in c4b up2 r15 r2 r3 thumbs; do AB=default python tools/time_config.py $c; done
  for t in 96 128; do AB=tile${t}x64 SLIMG_B200_TILE_W=$t python tools/time_config.py c4b; done
  AB=tile64x32 SLIMG_B200_TILE_W=64 SLIMG_B200_TILE_H=32 python tools/time_config.py r2
  AB=tile64x64 SLIMG_B200_TILE_W=64 SLIMG_B200_TILE_H=64 python tools/time_config.py r3
) > gpurun_out/configs_e.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_paths_jpeg.py -m gpu -q -x > gpurun_out/pytest_gpu_e.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_e.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_tile -s 1 -c 1 -o gpurun_out/tile_c4b_full2 python tools/time_config.py c4b > gpurun_out/ncu_tile.log 2>&1
