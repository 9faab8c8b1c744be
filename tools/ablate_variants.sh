#!/bin/bash
# Tuning aid: tools/ablate_annotate.py for each prebuilt variant library.
for v in "" "$@"; do
  if [ -n "$v" ]; then export PCS_B200_LIB=$PWD/point_cloud_segmentation_b200/csrc/build/variants/$v.so; fi
  echo "== ${v:-default}"; python tools/ablate_annotate.py | grep -E "map\+ann\+mask|ann only"
done
