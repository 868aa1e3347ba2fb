set -x
mkdir -p gpurun_out
( for c in c3 c3conv c4a c1 r6 r55 r3 r2 r15 thumb thumbs; do AB=default timeout 120 python tools/time_config.py $c; done
  for c in c3 c4a c1 r3 r2; do AB=cohorts1 SLIMG_B200_COHORTS=1 timeout 120 python tools/time_config.py $c; done
) > gpurun_out/configs_m.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_m.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_m.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_m.json 2> gpurun_out/bench_m.err; echo "bench rc=$?" >> gpurun_out/bench_m.err
