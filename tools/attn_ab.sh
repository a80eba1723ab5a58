#!/bin/bash
# A/B of attention builds on the GPU box: default library + every tools/ab/attn_*.so, three bench shapes each.
for lib in "" tools/ab/attn_*.so; do
  for shape in ${SHAPES:-cfg2 cfg3 cfg1}; do
    echo -n "${lib:-default} $shape: "
    MM_B200_LIB=$lib timeout 120 python tools/attn_bench.py --shape $shape --iters 20 2>&1 | tail -1
  done
done
