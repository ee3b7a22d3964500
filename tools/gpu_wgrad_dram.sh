#!/bin/bash
# DRAM bytes of one k_wgrad2 launch at bench size for the library in DSDF_LIB:  tools/gpu_wgrad_dram.sh <lib.so>
D=$PWD/medical-shape-disentanglement_b200/dsdf_b200
DSDF_LIB=$D/$1 NCU_B=64 timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,launch__grid_size \
  --clock-control none -k regex:k_wgrad2 -s 2 -c 2 --csv python tools/ncu_target.py 2>/dev/null | grep -E "k_wgrad2" | awk -F'","' '{print $(NF-2), $(NF-1), $NF}'
