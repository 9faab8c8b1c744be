#!/bin/bash
# Tuning aid: runs the GPU parity tests touched by the hot kernel, then bench.py for each prebuilt variant library
# (point_cloud_segmentation_b200/_build.py build_variant) selected through PCS_B200_LIB.  Usage: sweep_variants.sh [names...]
python -m pytest tests/test_gpu_annotate.py tests/test_gpu_map.py tests/test_gpu_mesh.py -m gpu -x -q 2>&1 | tail -3
for v in "" "$@"; do
  if [ -n "$v" ]; then export PCS_B200_LIB=$PWD/point_cloud_segmentation_b200/csrc/build/variants/$v.so; fi
  python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-mesh | python -c "
import json,sys
j=json.loads(sys.stdin.read()); print('${v:-default}', 'value %.1f G/s'%(j['value']/1e9), 'kernel_ms %.3f'%j['roofline']['kernel_ms'], 'frac %.3f'%j['roofline']['frac'])"
done
