set -x
mkdir -p gpurun_out
( for c in odd24 odd2 c2; do AB=default timeout 120 python tools/time_config.py $c; done ) > gpurun_out/configs_af.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_af.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_af.log
