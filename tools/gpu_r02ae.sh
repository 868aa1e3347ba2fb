set -x
mkdir -p gpurun_out
( for c in odd24 odd2 c3conv; do AB=default timeout 120 python tools/time_config.py $c; done
  AB=norepitch SLIMG_B200_NO_REPITCH=1 timeout 120 python tools/time_config.py odd2 ) > gpurun_out/configs_ae.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_ae.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_ae.log
grep -c generic gpurun_out/parity_report.txt >> gpurun_out/pytest_gpu_ae.log
