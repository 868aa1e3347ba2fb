set -x
mkdir -p gpurun_out
( for c in r2 r15 r3 c4a; do
    AB=default timeout 120 python tools/time_config.py $c
    AB=pf0 SLIMG_B200_PREFETCH_STEPS=0 timeout 120 python tools/time_config.py $c
    AB=pf4 SLIMG_B200_PREFETCH_STEPS=4 timeout 120 python tools/time_config.py $c
    AB=pf0_dbg3 SLIMG_B200_PREFETCH_STEPS=0 SLIMG_B200_VF_DEBUG=3 timeout 120 python tools/time_config.py $c
    AB=co1 SLIMG_B200_COHORTS=1 timeout 120 python tools/time_config.py $c
  done
) > gpurun_out/configs_w.txt 2>&1
