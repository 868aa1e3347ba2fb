set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 20 > gpurun_out/bench_aq.json 2> gpurun_out/bench_aq.err; echo "rc=$?" >> gpurun_out/bench_aq.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_vfirst -s 1 -c 1 -o gpurun_out/vfirst_full python bench.py --batch 16 --steps 1 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_tile -s 1 -c 1 -o gpurun_out/tile_full python tools/time_config.py c4b > gpurun_out/ncu_tile.log 2>&1
( for c in c1 c2 c3 c3conv c4a c4b up2 r15 r2 r3 r55 r6 thumb thumb24 thumbs; do AB=default python tools/time_config.py $c; done ) > gpurun_out/configs_u.txt 2>&1
