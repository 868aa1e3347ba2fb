mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
export SLIMG_B200_DEBUG=1
bash tools/ab.sh "$@"
grep slimg_b200 gpurun_out/ab.err | sort | uniq -c
