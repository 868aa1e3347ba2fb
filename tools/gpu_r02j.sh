set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_planes.py tests/test_gpu_paths_jpeg.py tests/test_gpu_uniffi.py tests/test_reference_binding.py -m gpu -q > gpurun_out/pytest_gpu_j.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_j.log
timeout 900 python bench.py --steps 20 --no-e2e --no-cpu > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err; echo "rc=$?" >> gpurun_out/bench_j.err
