set -x
mkdir -p gpurun_out
L=$PWD/slimg_b200/lib
( for rep in 1 2; do for c in c3 r2 c4a; do
    AB=new timeout 120 python tools/time_config.py $c
    AB=old SLIMG_B200_LIB=$L/libslimg_b200_old.so timeout 120 python tools/time_config.py $c
    AB=oldsig SLIMG_B200_LIB=$L/libslimg_b200_oldsig.so timeout 120 python tools/time_config.py $c
    AB=oldprod SLIMG_B200_LIB=$L/libslimg_b200_oldprod.so timeout 120 python tools/time_config.py $c
  done; done
) > gpurun_out/configs_y.txt 2>&1
