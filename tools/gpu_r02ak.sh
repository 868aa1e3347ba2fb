set -x
mkdir -p gpurun_out
timeout 600 python tools/ratio_scan.py 6000 4000 16 > gpurun_out/ratio_scan_24mp.txt 2>&1
timeout 600 python tools/ratio_scan.py 4000 3000 32 > gpurun_out/ratio_scan_12mp.txt 2>&1
