#!/bin/bash
# ncu record of the HBM-bound kernels (VERDICT r2 missing #3): duration + DRAM bytes per launch, clocks untouched.
# Each target is run once WITHOUT ncu first (it must exit 0), then under ncu.  Output: gpurun_out/r3_ncu_*.csv,
# summarised by tools/sweeps/r3_ncu_table.py into profiles/r3_ncu_hbm_kernels.md.
set -u
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum"
K='regex:layernorm|embed|gather|scatter|segcopy|attention'
run() {  # name, command...
  local name=$1; shift
  "$@" > gpurun_out/r3_ncu_$name.plain.log 2>&1 || { echo "$name: target failed without ncu"; return; }
  ncu --metrics $M --clock-control none -k "$K" --csv --log-file gpurun_out/r3_ncu_$name.csv "$@" > gpurun_out/r3_ncu_$name.log 2>&1
  echo "$name: ncu rc=$? rows=$(wc -l < gpurun_out/r3_ncu_$name.csv)"
}
run b32 python tools/ncu_target.py 32 128
run b256 python tools/ncu_target.py 256 128
run requests python tools/sweeps/r3_ncu_requests_target.py
