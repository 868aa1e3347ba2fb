set -x
mkdir -p gpurun_out
( for c in odd24 odd2 c3 c4a; do AB=default python tools/time_config.py $c; done ) > gpurun_out/configs_as.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_u.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_u.log
timeout 900 python bench.py --steps 20 > gpurun_out/bench_chain.json 2> gpurun_out/bench_chain.err; echo "rc=$?" >> gpurun_out/bench_chain.err
timeout 600 python bench.py --steps 100 --no-configs --no-cpu > gpurun_out/bench_chain_100.json 2>> gpurun_out/bench_chain.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_vfirst -s 1 -c 1 -o gpurun_out/vfirst_full python bench.py --batch 16 --steps 1 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_full.log 2>&1
