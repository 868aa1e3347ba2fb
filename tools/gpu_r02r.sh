set -x
mkdir -p gpurun_out
( AB=default timeout 120 python tools/time_config.py c4a
  for v in 2 3; do AB=vbufs$v SLIMG_B200_VBUFS=$v timeout 120 python tools/time_config.py c4a; done
  for sc in 40 48 64; do AB=cols$sc SLIMG_B200_STRIP_COLS=$sc timeout 120 python tools/time_config.py c4a; done
  for sc in 40 48 64; do AB=cols${sc}_vbufs3 SLIMG_B200_VBUFS=3 SLIMG_B200_STRIP_COLS=$sc timeout 120 python tools/time_config.py c4a; done
  AB=default timeout 120 python tools/time_config.py c1
  for v in 2 4; do AB=vbufs$v SLIMG_B200_VBUFS=$v timeout 120 python tools/time_config.py c1; done
  for sc in 32 48; do AB=cols$sc SLIMG_B200_STRIP_COLS=$sc timeout 120 python tools/time_config.py c1; done
  AB=default timeout 120 python tools/time_config.py r3
  for v in 3; do AB=vbufs$v SLIMG_B200_VBUFS=$v timeout 120 python tools/time_config.py r3; AB=vbufs$v SLIMG_B200_VBUFS=$v timeout 120 python tools/time_config.py r2; done
  AB=default timeout 120 python tools/time_config.py r2
  for c in c3 c4b up2 c2; do AB=default timeout 120 python tools/time_config.py $c; done
) > gpurun_out/configs_r.txt 2>&1
