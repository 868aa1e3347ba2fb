set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_an.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_an.log
timeout 600 python bench.py --steps 20 --no-pageable > gpurun_out/bench_an.json 2> gpurun_out/bench_an.err; echo "rc=$?" >> gpurun_out/bench_an.err
