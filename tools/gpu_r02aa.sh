set -x
mkdir -p gpurun_out
L=$PWD/slimg_b200/lib
( for rep in 1 2; do for c in c4b up2; do
    AB=minb3 timeout 120 python tools/time_config.py $c
    AB=minb4 SLIMG_B200_LIB=$L/libslimg_b200_tile4.so timeout 120 python tools/time_config.py $c
  done; done
  for c in r15 r2; do
    AB=tile_minb3 SLIMG_B200_RESIZE_PATH=tile timeout 120 python tools/time_config.py $c
    AB=tile_minb4 SLIMG_B200_RESIZE_PATH=tile SLIMG_B200_LIB=$L/libslimg_b200_tile4.so timeout 120 python tools/time_config.py $c
  done
) > gpurun_out/configs_aa.txt 2>&1
