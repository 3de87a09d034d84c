#!/bin/bash
# Rows per warp of k_burr_chain_cols at batch 64: ncu durations of the chain kernel and the hysteresis for a few settings.
TAG=${1:-chain}
mkdir -p gpurun_out
for R in 0 4 6 12 16; do
  UB_CHAIN_ROWS=$R timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_burr_chain_cols|k_hyst_frame|k_masks' -c 12 --csv \
      --log-file gpurun_out/${TAG}_R$R.csv python tools/profile_step.py 3 64 > gpurun_out/${TAG}_R$R.log 2>&1
  echo "R=$R: $(grep -E 'k_burr_chain|k_hyst|k_masks' gpurun_out/${TAG}_R$R.csv | awk -F'","' '{print $5, $NF}' | sed 's/(.*) / /' | tr '\n' ';' | cut -c1-400)"
done
