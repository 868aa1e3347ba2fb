# round-2 GPU pass B: the integer tile kernel -- parity (full suite), then time against the banded pipelines / f32 kernels
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_b.log
( for c in c4b up2 r15 r2 r3 thumbs; do AB=default python tools/time_config.py $c; done
  for c in r15 r2 r3; do AB=notile SLIMG_B200_RESIZE_PATH=notile python tools/time_config.py $c; done
  for c in c4b up2 thumbs; do AB=generic SLIMG_B200_RESIZE_PATH=generic python tools/time_config.py $c; done
  for c in c1 c4a; do AB=tile SLIMG_B200_RESIZE_PATH=tile python tools/time_config.py $c; done
  for c in c1 c4a thumb thumb24; do AB=default python tools/time_config.py $c; done
  for t in 16 32 48 64; do AB=tile${t} SLIMG_B200_TILE_W=$t SLIMG_B200_TILE_H=$t python tools/time_config.py c4b; AB=tile${t} SLIMG_B200_TILE_W=$t SLIMG_B200_TILE_H=$t python tools/time_config.py r2; done
  AB=tile64x32 SLIMG_B200_TILE_W=64 SLIMG_B200_TILE_H=32 python tools/time_config.py r2
  AB=tile64x32 SLIMG_B200_TILE_W=64 SLIMG_B200_TILE_H=32 python tools/time_config.py c4b
) > gpurun_out/configs_b.txt 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_tile -s 1 -c 1 -o gpurun_out/tile_c4b_full python tools/time_config.py c4b > gpurun_out/ncu_tile.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_tile -s 1 -c 1 -o gpurun_out/tile_r2_full python tools/time_config.py r2 >> gpurun_out/ncu_tile.log 2>&1
ls -la gpurun_out
