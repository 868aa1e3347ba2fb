# round-1d GPU pass: tests, smoke, bench (both orders, reference arm), other configs, ncu launch list + full captures
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt
nproc >> gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_chain.json 2> gpurun_out/bench_chain.err; echo "rc=$?" >> gpurun_out/bench_chain.err
timeout 600 python bench.py --steps 20 --no-cpu --no-e2e > gpurun_out/bench_chain_burst.json 2>> gpurun_out/bench_chain.err
timeout 600 python bench.py --order convert --no-cpu > gpurun_out/bench_convert.json 2> gpurun_out/bench_convert.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
bash tools/ab_cfg.sh "c1 c2 c3 c3conv c4a c4b" "AB=default"
python tools/time_jpeg.py > gpurun_out/time_jpeg.txt 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:resize_vfirst -s 1 -c 1 -o gpurun_out/vfirst_full python bench.py --batch 16 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jpeg_frontend -s 1 -c 1 -o gpurun_out/k3_full python tools/time_jpeg.py > gpurun_out/ncu_k3.log 2>&1
ls -la gpurun_out
