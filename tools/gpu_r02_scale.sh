# usage: gpu_r02_scale.sh N   (under gpurun --gpus N)
set -x
N=$1
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/smi_${N}gpu.txt; nproc >> gpurun_out/smi_${N}gpu.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "rc=$?" >> gpurun_out/bench_${N}gpu.err
timeout 600 python -m pytest tests/test_gpu_paths_jpeg.py -m gpu -q -k "multi_device or mixed or batch" > gpurun_out/pytest_${N}gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_${N}gpu.log
