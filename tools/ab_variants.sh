#!/bin/bash
# A/B of kernel build variants (build/lib_*.so): per-kernel times of one bench step
for lib in "" build/lib_*.so; do
  echo "== ${lib:-default}"
  HIST_B200_LIB=${lib:+$PWD/$lib} bash tools/launch_times.sh | grep -E "absorb_k|build|stream|count_cells"
done
