set -x
mkdir -p gpurun_out
( for c in odd24 c3conv odd2 odd2a; do AB=default timeout 120 python tools/time_config.py $c; done ) > gpurun_out/configs_ad.txt 2>&1
