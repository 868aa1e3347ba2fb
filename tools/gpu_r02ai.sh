set -x
mkdir -p gpurun_out
( for c in r748 r8; do
    for v in 2 3 4; do for e in 0 1; do
      AB=cols16_vb${v}_early$e SLIMG_B200_STRIP_COLS=16 SLIMG_B200_VBUFS=$v SLIMG_B200_EARLY_GO=$e SLIMG_B200_DEBUG=1 timeout 120 python tools/time_config.py $c 2>&1 | grep "vplan\|us/image"
    done; done
  done
  for v in 2 3; do AB=c3_vb$v SLIMG_B200_VBUFS=$v SLIMG_B200_DEBUG=1 timeout 120 python tools/time_config.py c3conv 2>&1 | grep "vplan\|us/image"; done
) > gpurun_out/configs_ai.txt 2>&1
