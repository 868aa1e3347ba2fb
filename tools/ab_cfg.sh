# usage: tools/ab_cfg.sh "cfg cfg ..." "ENV=VAL ..." ...  -- time_config.py per config and variant
mkdir -p gpurun_out; : > gpurun_out/abcfg.txt
cfgs="$1"; shift
for c in $cfgs; do for v in "$@"; do env AB="$v" $v timeout 120 python tools/time_config.py $c >> gpurun_out/abcfg.txt 2>> gpurun_out/abcfg.err; done; done
cat gpurun_out/abcfg.txt
