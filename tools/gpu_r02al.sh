set -x
mkdir -p gpurun_out
( for c in c3 c3conv c1 c4a r2 thumb; do AB=final timeout 120 python tools/time_config.py $c; done ) > gpurun_out/configs_al.txt 2>&1
for c in c3 c4a r2; do SLIMG_B200_LIB=$PWD/slimg_b200/lib/libslimg_b200_exp.so SLIMG_B200_VF_DEBUG=32 timeout 120 python tools/time_config.py $c > gpurun_out/role_waits_$c.txt 2>&1; done
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_al.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_al.log
timeout 600 python bench.py --steps 20 > gpurun_out/bench_al.json 2> gpurun_out/bench_al.err; echo "rc=$?" >> gpurun_out/bench_al.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_vfirst -s 1 -c 1 -o gpurun_out/vfirst_full python bench.py --batch 16 --steps 1 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_full.log 2>&1
