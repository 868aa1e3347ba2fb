# usage: tools/sweep_c1.sh "cfg ..." "vbufs ..." "cols ..."   -- V-buffer count x strip width sweep with the plan each combination gets
mkdir -p gpurun_out; : > gpurun_out/sweep.txt
for cfg in ${1:-c1 c4a}; do
for vb in ${2:-2 3 4}; do for cols in ${3:-0 24 32 40 48 56 64}; do
  if [ $cols = 0 ]; then unset SLIMG_B200_STRIP_COLS; else export SLIMG_B200_STRIP_COLS=$cols; fi
  AB="vb=$vb cols=$cols" SLIMG_B200_VBUFS=$vb SLIMG_B200_DEBUG=1 timeout 120 python tools/time_config.py $cfg 2>&1 | grep -E "vplan|us/image" | sed 's/\[slimg_b200\] vplan//' | tr '\n' ' ' >> gpurun_out/sweep.txt; echo >> gpurun_out/sweep.txt
done; done; done
cat gpurun_out/sweep.txt
