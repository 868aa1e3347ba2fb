set -x
mkdir -p gpurun_out
L=$PWD/slimg_b200/lib
( for c in c4b up2; do
    AB=full timeout 120 python tools/time_config.py $c
    for v in TILE_SKIP_P1 TILE_SKIP_P2 TILE_SKIP_ST; do AB=$v SLIMG_B200_LIB=$L/libslimg_b200_$v.so timeout 120 python tools/time_config.py $c; done
  done
) > gpurun_out/configs_ab.txt 2>&1
