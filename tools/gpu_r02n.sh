set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_n.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_n.log
( for c in c3 c1 thumb r6 r55; do AB=default timeout 120 python tools/time_config.py $c; done
  for c in c1 thumb r55; do AB=cohorts2 SLIMG_B200_COHORTS=2 timeout 120 python tools/time_config.py $c; done
) > gpurun_out/configs_n.txt 2>&1
