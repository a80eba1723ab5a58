"""Standalone splice passes at BASELINE sizes, for the HBM roofline and its ncu traffic record.

  scatter      cfg2: 256 images x 576 tokens x 4096 bf16 spliced in place into [256, 640, 4096]   (2.416 GB algorithmic)
  embed+splice cfg4: inputs_embeds [64, 8192, 4096] bf16 produced from ids, table and the projected stack in one pass
               (mm_embed_splice_fused; 2 x 64 x 8192 x 4096 x 2 B = 8.59 GB algorithmic)

Plain run: CUDA-event timings -> one JSON line.  Under ncu (tools/ncu_round.sh): the launches to capture are
`move_rows_kernel` #3.. (scatter; launch 1 is the embed gather, 2 a warm-up) and the last ones (fused).
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multimeditron_b200 import ops, synth  # noqa: E402

REPS = int(os.environ.get("SPLICE_REPS", "10"))
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)


def timed(fn, reps):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(reps):
        fn()
    ev[1].record()
    torch.cuda.synchronize()
    return ev[0].elapsed_time(ev[1]) / reps


out = {}
# ---- cfg2 scatter ------------------------------------------------------------------------------------------
B, P, H, V, n_text = 256, 576, 4096, 128_256, 64
att = V - 4
ids, _ = synth.make_sequences(B, P + n_text, 1, P, V - 4, att, V - 3, V - 2, V - 1, seed=1)
ids_d = ids.to(dev)
table = (torch.randn(V, H, device=dev) * 0.02).bfloat16()
proj = torch.randn(B * P, H, device=dev).bfloat16()
dest_rows, _, _ = ops.attach_positions(ids_d, att, expected_rows=B * P, want_src_of_pos=False)
emb = ops.embed_gather(table, ids_d)                                  # move_rows launch 1 (MODE 3)
ops.splice_scatter(emb.view(-1, H), proj, dest_rows)                  # launch 2 (warm-up)
ms = timed(lambda: ops.splice_scatter(emb.view(-1, H), proj, dest_rows), REPS)
nbytes = 2.0 * B * P * H * 2
out["scatter_cfg2"] = {"bytes": nbytes, "ms": ms, "GBps": nbytes / ms / 1e6}
del emb, proj, ids_d
# ---- cfg4 fused embed + splice ----------------------------------------------------------------------------------
Bs, S, per_row, P4 = 64, 8192, 4, 576
ids4, _ = synth.make_sequences(Bs, S, per_row, P4, V - 4, att, V - 3, V - 2, V - 1, seed=2)
ids4_d = ids4.to(dev)
n_rows = Bs * per_row * P4
src = torch.randn(n_rows, H, device=dev).bfloat16()
_, sop, _ = ops.attach_positions(ids4_d, att, expected_rows=n_rows, want_src_of_pos=True)
o4 = torch.empty(Bs, S, H, dtype=torch.bfloat16, device=dev)
ops.embed_splice_fused(table, ids4_d, sop, src, out=o4)
ms4 = timed(lambda: ops.embed_splice_fused(table, ids4_d, sop, src, out=o4), max(3, REPS // 2))
nb4 = 2.0 * Bs * S * H * 2
out["embed_splice_cfg4"] = {"bytes": nb4, "ms": ms4, "GBps": nb4 / ms4 / 1e6}
print(json.dumps(out))
