#!/bin/bash
# Static SASS evidence per hot kernel: instruction histogram (top 25) and the lines that prove the Blackwell paths
# (UBLKCP = cp.async.bulk, SYNCS = mbarrier, UTCHMMA / LDTM = tcgen05, packed FFMA2 / FADD2).  usage: bash tools/sass_excerpt.sh > profiles/<tag>_sass.txt
LIB=mmwave_radar_processing_b200/lib/libradar_b200.so
for K in 'rd_reg_kernel.*Li63ELi70ELi12ELi2ELi1ELi3ELi1ELi1' 'rd_reg_kernel.*Li100ELi30' 'rd_reg_kernel.*Li90ELi100' 'cfar_kernelENS' 'range_angle_kernel' 'point_cloud_thread_kernel' 'sar_tc_kernelILi2ELi1ELi16ELb1' 'sar_range_fft_reg_kernel' 'dbs_combine_kernel' 'dbs_kernel'; do
  SYM=$(cuobjdump -elf $LIB 2>/dev/null | grep -oE "_ZN2rb[A-Za-z0-9_]*" | grep -E "$K" | sort -u | head -1)
  [ -z "$SYM" ] && { echo "== $K: symbol not found"; continue; }
  echo "== $SYM"
  cuobjdump -sass -fun "$SYM" $LIB > /tmp/sass_one.txt 2>/dev/null
  echo "instructions: $(grep -cE '^\s+/\*[0-9a-f]{4,5}\*/' /tmp/sass_one.txt)"
  grep -E '^\s+/\*[0-9a-f]{4,5}\*/' /tmp/sass_one.txt | awk '{ if ($2 ~ /^@/) print $3; else print $2}' | sed 's/;//' | sort | uniq -c | sort -rn | head -25
  echo "-- Blackwell-specific lines (first of each kind)"
  grep -E "UBLKCP|SYNCS|UTCHMMA|LDTM|UTCBAR|REDG|F2FP|CCTL|FFMA2|FADD2|I2F.S16|STS.128|STG.E.128" /tmp/sass_one.txt | sed 's/ *\/\* 0x[0-9a-f]* \*\///' | awk '{k=$2; if (k ~ /^@/) k=$3; sub(/\..*/,"",k); if (!seen[k]++) print}' | head -12
  echo
done
