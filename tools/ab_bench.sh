#!/bin/bash
# Same-box A/B of writer variants: tools/ab_bench.sh <libA.so> <libB.so> ...  (each run twice, interleaved)
SO=kubernetes-gpu-exporter_b200/csrc/libkgx.so
cp $SO /tmp/libkgx_orig.so
for round in 1 2; do
  for lib in "$@"; do
    cp "$lib" $SO
    timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); k=d['kernel_ms']; print('%-28s step %.3f ms  ingest %.4f  render %.4f' % ('$lib'.split('/')[-1], d['ms_per_step'], k['ingest'], k['render']))"
  done
done
cp /tmp/libkgx_orig.so $SO
