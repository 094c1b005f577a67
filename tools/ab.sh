#!/bin/bash
# Same-box A/B of library builds: tools/ab.sh <lib.so> ...   (each run twice, interleaved; KGX_LIB selects the build)
for round in 1 2; do
  for lib in "$@"; do
    KGX_LIB=$PWD/$lib timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); k=d['kernel_ms']; print('%-28s step %.3f ms  ingest %.4f  render %.4f other %.4f' % ('$lib'.split('/')[-1], d['ms_per_step'], k['ingest'], k['render'], k['other']))"
  done
done
