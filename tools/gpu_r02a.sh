# round-2 GPU pass A: tests, smoke, bench (all configs), tc vs vfirst, K3 stage-1 prefetch variants, ncu launch list + full captures
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt
nproc >> gpurun_out/smi.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 > gpurun_out/bench_chain.json 2> gpurun_out/bench_chain.err; echo "rc=$?" >> gpurun_out/bench_chain.err
SLIMG_B200_RESIZE_PATH=tc timeout 300 python bench.py --steps 20 --no-cpu --no-e2e --no-configs > gpurun_out/bench_tc.json 2> gpurun_out/bench_tc.err
for v in 0 1 2 3 4; do SLIMG_B200_JPEG_VAR=$v python tools/time_jpeg.py; done > gpurun_out/time_jpeg.txt 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_vfirst -s 1 -c 1 -o gpurun_out/vfirst_full python bench.py --batch 16 --steps 1 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_full.log 2>&1
SLIMG_B200_RESIZE_PATH=tc timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_tc -s 1 -c 1 -o gpurun_out/tc_full python bench.py --batch 16 --steps 1 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_tc.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jpeg_frontend -s 1 -c 1 -o gpurun_out/k3_full python tools/time_jpeg.py > gpurun_out/ncu_k3.log 2>&1
ls -la gpurun_out
