set -x
mkdir -p gpurun_out
( for c in r748 r7 r8 odd24; do
    AB=default timeout 120 python tools/time_config.py $c
    AB=cols16 SLIMG_B200_STRIP_COLS=16 timeout 120 python tools/time_config.py $c
  done ) > gpurun_out/configs_ah.txt 2>&1
