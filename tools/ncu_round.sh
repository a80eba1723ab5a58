#!/bin/bash
# usage (under gpurun): bash tools/ncu_round.sh <tag>
# Launch list (small batch, shares only) + one `--set full` capture of the top kernels at the bench's full size.
tag=${1:-r01}
mkdir -p gpurun_out
SMALL="python bench.py --steps 1 --warmup 3 --images-per-gpu 32 --no-cpu-baseline"
FULL="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$SMALL > gpurun_out/plain_small.log 2>&1 &&
# only this library's kernels are counted (weight packing at load time launches thousands of torch kernels)
ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:gemm_bf16_tn_kernel|attention3_fwd|attention_fwd|layernorm_kernel|move_rows|attach_|cls_rows|patch_rows|resize_|weighted_expert" -s 396 -c 140 --csv \
    --log-file gpurun_out/${tag}_launches.csv $SMALL > gpurun_out/ncu_small.log 2>&1
echo "launch list exit $?"
$FULL > gpurun_out/plain_full.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_tn_kernel -s 320 -c 6 \
    -o gpurun_out/${tag}_gemm $FULL > gpurun_out/ncu_gemm.log 2>&1
echo "gemm capture exit $?"
ncu --set full --clock-control none --import-source on -k "regex:attention3_fwd|attention_fwd" -s 60 -c 2 \
    -o gpurun_out/${tag}_attn $FULL > gpurun_out/ncu_attn.log 2>&1
echo "attn capture exit $?"
# HBM-bound splice kernels: a dedicated target (the step itself fuses the splice into the projector GEMM's epilogue), so the
# capture cannot land on load-time launches of another kernel: launch 1 = embed gather, 2 = scatter warm-up, 3-4 = scatter
# (move_rows_kernel<0>, cfg2 shape), 5-7 = fused embed+splice (move_rows_kernel<2>, cfg4 shape)
python tools/splice_roofline.py > gpurun_out/${tag}_splice_plain.json 2> gpurun_out/splice_plain.err &&
SPLICE_REPS=2 ncu --set full --clock-control none --import-source on -k regex:move_rows_kernel -s 2 -c 5 \
    -o gpurun_out/${tag}_splice python tools/splice_roofline.py > gpurun_out/ncu_splice.log 2>&1
echo "splice capture exit $?"
ls -la gpurun_out
