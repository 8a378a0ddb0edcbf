#!/bin/bash
# Sweeps the attention kernel's switches on the UNet's shapes (tools/attn_case.py); run under gpurun.
out=${1:-gpurun_out/attn_sweep.txt}
: > $out
for layout in 0 1; do
  for pb in 0 1; do
    [ $layout = 1 ] && [ $pb = 1 ] && continue
    for emu in 0 2 3 1; do
      echo "== layout=$layout pair_bar=$pb emu=$emu" >> $out
      B200SD_ATTN_LAYOUT=$layout B200SD_ATTN_PAIR_BAR=$pb B200SD_ATTN_EXP_EMU=$emu python tools/attn_case.py 16 5 >> $out 2>&1
    done
  done
done
