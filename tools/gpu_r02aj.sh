set -x
mkdir -p gpurun_out
( for c in r748 r8 r8b r86 r7 odd24 c3 c3conv c1 c4a thumb; do AB=default timeout 120 python tools/time_config.py $c; done
  for c in r86; do AB=cols16_vb2 SLIMG_B200_STRIP_COLS=16 SLIMG_B200_VBUFS=2 timeout 120 python tools/time_config.py $c; done
) > gpurun_out/configs_aj.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_aj.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_aj.log
