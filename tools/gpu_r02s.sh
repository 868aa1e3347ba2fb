set -x
mkdir -p gpurun_out
( for c in c4a c1 r6 r55 c3 thumb; do AB=default timeout 120 python tools/time_config.py $c; done
  for v in 2 4; do AB=vbufs$v SLIMG_B200_VBUFS=$v timeout 120 python tools/time_config.py c1; done
  for v in 2 3; do AB=vbufs$v SLIMG_B200_VBUFS=$v timeout 120 python tools/time_config.py c4a; done
  for v in 2 4; do AB=vbufs$v SLIMG_B200_VBUFS=$v timeout 120 python tools/time_config.py r6; done
  AB=cols24_co1 SLIMG_B200_STRIP_COLS=24 timeout 120 python tools/time_config.py c3
  AB=cols24_co2 SLIMG_B200_STRIP_COLS=24 SLIMG_B200_COHORTS=2 timeout 120 python tools/time_config.py c3
  AB=cols24_co2_vbufs3 SLIMG_B200_STRIP_COLS=24 SLIMG_B200_COHORTS=2 SLIMG_B200_VBUFS=3 timeout 120 python tools/time_config.py c3
  AB=cols32_co2_vbufs3 SLIMG_B200_STRIP_COLS=32 SLIMG_B200_COHORTS=2 SLIMG_B200_VBUFS=3 timeout 120 python tools/time_config.py c3
  AB=vbufs3 SLIMG_B200_VBUFS=3 timeout 120 python tools/time_config.py c3
  AB=co3 SLIMG_B200_COHORTS=3 timeout 120 python tools/time_config.py c3
  AB=unit3 SLIMG_B200_UNIT=3 timeout 120 python tools/time_config.py c3
  AB=unit5 SLIMG_B200_UNIT=5 timeout 120 python tools/time_config.py c3
) > gpurun_out/configs_s.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_s.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_s.log
