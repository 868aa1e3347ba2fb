set -x
mkdir -p gpurun_out
( for c in c3 c3conv; do for co in 1 2 3; do for u in 4 5; do
    AB=co${co}_u$u SLIMG_B200_COHORTS=$co SLIMG_B200_UNIT=$u timeout 120 python tools/time_config.py $c
  done; done; done
  for c in c4a c1 r6 r55 r3 thumb; do for co in 2; do for u in 3 4 5; do
    AB=co${co}_u$u SLIMG_B200_COHORTS=$co SLIMG_B200_UNIT=$u timeout 120 python tools/time_config.py $c
  done; done; done
  for c in c4a r3; do for u in 3 4 5; do AB=co3_u$u SLIMG_B200_COHORTS=3 SLIMG_B200_UNIT=$u timeout 120 python tools/time_config.py $c; done; done
  AB=co3_u5_vbufs3 SLIMG_B200_COHORTS=3 SLIMG_B200_UNIT=5 SLIMG_B200_VBUFS=3 timeout 120 python tools/time_config.py c3
  AB=co2_u5_vbufs3 SLIMG_B200_COHORTS=2 SLIMG_B200_UNIT=5 SLIMG_B200_VBUFS=3 timeout 120 python tools/time_config.py c3
) > gpurun_out/configs_t.txt 2>&1
