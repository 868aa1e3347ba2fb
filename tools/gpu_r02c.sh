# round-2 GPU pass C: tile kernel v2 (integer epilogue, packed limbs), tcgen05 kernel v2 (8-row blocks, 4 accumulator buffers, L2 prefetch, fill CTAs)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_c.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_c.log
( for c in c4b up2 r15 r2 r3 thumbs; do AB=default python tools/time_config.py $c; done
  AB=nopack SLIMG_B200_TILE_NOPACK=1 python tools/time_config.py c4b
  for t in 32 48; do AB=tile${t} SLIMG_B200_TILE_W=$t SLIMG_B200_TILE_H=$t python tools/time_config.py c4b; done
  AB=tile64x32 SLIMG_B200_TILE_W=64 SLIMG_B200_TILE_H=32 python tools/time_config.py r2
  AB=tile48 SLIMG_B200_TILE_W=48 SLIMG_B200_TILE_H=48 python tools/time_config.py r2
  for c in c3 c1 c4a r3; do AB=tc SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py $c; done
  for c in c3 c1 c4a; do AB=default python tools/time_config.py $c; done
  for d in 1 2 8 9 11; do AB=tc_debug$d SLIMG_B200_VF_DEBUG=$d SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3; done
  for pf in 0 3 12; do AB=tc_pf$pf SLIMG_B200_TC_PREFETCH=$pf SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3; done
  for vb in 2 4; do AB=tc_vb$vb SLIMG_B200_TC_VBUFS=$vb SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3; done
  for sc in 16 32; do AB=tc_cols$sc SLIMG_B200_STRIP_COLS=$sc SLIMG_B200_RESIZE_PATH=tc python tools/time_config.py c3; done
) > gpurun_out/configs_c.txt 2>&1
SLIMG_B200_RESIZE_PATH=tc timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_tc -s 1 -c 1 -o gpurun_out/tc2_full python bench.py --batch 16 --steps 1 --warmup 3 --no-e2e --no-cpu --no-configs > gpurun_out/ncu_tc2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:resize_tile -s 1 -c 1 -o gpurun_out/tile_c4b_full python tools/time_config.py c4b > gpurun_out/ncu_tile.log 2>&1
ls -la gpurun_out
