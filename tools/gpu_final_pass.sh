# round-2 GPU final pass: full suite, smoke, bench (burst + sustained), launch list, full captures of the default kernels
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt; nproc >> gpurun_out/smi.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_u.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_i.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 > gpurun_out/bench_chain.json 2> gpurun_out/bench_chain.err; echo "rc=$?" >> gpurun_out/bench_chain.err
timeout 600 python bench.py --steps 100 --no-configs --no-cpu > gpurun_out/bench_chain_100.json 2>> gpurun_out/bench_chain.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_vfirst -s 1 -c 1 -o gpurun_out/vfirst_full python bench.py --batch 16 --steps 1 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jpeg_frontend -s 1 -c 1 -o gpurun_out/k3_full python tools/time_jpeg.py > gpurun_out/ncu_k3.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_tile -s 1 -c 1 -o gpurun_out/tile_full python tools/time_config.py c4b > gpurun_out/ncu_tile.log 2>&1
( for c in c1 c2 c3 c3conv c4a c4b up2 r15 r2 r3 r55 r6 thumb thumb24 thumbs; do AB=default python tools/time_config.py $c; done ) > gpurun_out/configs_u.txt 2>&1
ls -la gpurun_out
