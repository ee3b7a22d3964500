#!/bin/bash
# same-box A/B of two builds of the library: tools/gpu_ab.sh <libA.so> <libB.so> [eik]   (alternating, 3 rounds)
D=$PWD/medical-shape-disentanglement_b200/dsdf_b200
for i in 1 2 3; do
  for lib in "$1" "$2"; do
    DSDF_LIB=$D/$lib timeout 120 python tools/gpu_sections.py 20 $3 2>&1 | tail -1
  done
done
