set -x
mkdir -p gpurun_out
( for c in c4a c1 r6 r55 r3 r2 r15 thumbs c3; do
    AB=tpw_auto timeout 120 python tools/time_config.py $c
    AB=tpw1 SLIMG_B200_P2_TPW=1 timeout 120 python tools/time_config.py $c
  done
  for sc in 48 56; do AB=tpw_auto_cols$sc SLIMG_B200_STRIP_COLS=$sc timeout 120 python tools/time_config.py c4a; done
  for sc in 32 48; do AB=tpw_auto_cols$sc SLIMG_B200_STRIP_COLS=$sc timeout 120 python tools/time_config.py c1; done
  for v in 3; do AB=tpw_auto_vbufs$v SLIMG_B200_VBUFS=$v timeout 120 python tools/time_config.py c4a; done
) > gpurun_out/configs_q.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_q.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_q.log
