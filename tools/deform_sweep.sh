#!/bin/bash
# development: deformation launch pair under every accumulation mode, rows as built and tet-sorted, twice (alternating)
for rep in 1 2; do
  for layout in "" sorted; do
    for cfg in "auto 0" "ordered 0" "ordered 1" "padded 0" "plain 0"; do
      set -- $cfg
      echo -n "layout=${layout:-as_built} mode=$1 pitch4=$2 "
      D3GA_DEFORM_BACKWARD=$1 D3GA_DEFORM_PITCH4=$2 python tools/deform_probe.py metric $layout
    done
  done
done
