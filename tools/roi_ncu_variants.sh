#!/bin/bash
# DRAM / L2 traffic of the ROIAlign kernel under the development knobs (run on a B200 through gpurun).
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sector_hit_rate.pct,l1tex__m_xbar2l1tex_read_bytes.sum,smsp__inst_executed.sum
for v in "2 0" "0 0" "0 1" "2 1" "1 1"; do
  set -- $v
  echo "== PROMO=$1 ALIGN8=$2"
  OVM_ROI_SPLIT=400 OVM_ROI_PROMO=$1 OVM_ROI_ALIGN8=$2 ncu --metrics $M --clock-control none -k regex:roi_align_kernel --launch-skip 10 --launch-count 1 python tools/roi_bench.py --one 2>&1 | grep -E "^\s+(gpu__|dram__|lts__|l1tex__|smsp__)"
done
