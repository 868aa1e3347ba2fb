set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --no-configs --no-cpu > gpurun_out/bench_at.json 2> gpurun_out/bench_at.err; echo "rc=$?" >> gpurun_out/bench_at.err
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_at.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_at.log
