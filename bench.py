#!/usr/bin/env python
"""bench.py — image-embed tokens/s of the modality-embedding hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # B200 arm  (one process per GPU under torchrun for N>1)
    python bench.py --impl reference --gpus N --steps K ...   # reference arm: the CPU restatement on the host cores

Workload (config.workload): BASELINE.json configs[1] — CLIP ViT-L/14-336 + MLP projector into hidden 4096, bf16,
synthetic 336px images, one image per row of a 642-token sequence (576 attachment tokens + 2 delimiters + 64 text),
256 images per GPU (weak scaling: the per-GPU batch is fixed, rows are sharded by sample, no collective on the path).
A step = embed_modalities_with_text over one batch: embedding lookup, patch-embed, 24 encoder layers, projector,
splice.  `value` has inputs resident in HBM; `e2e` runs the same step through the public embedder call with HOST
(pinned) buffers, the H2D copies and a D2H read of the result inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC, UNIT = "image_embed_tokens_per_s", "tokens/s"
WORKLOADS = {
    # name: (tower, llm hidden, images/GPU, text tokens per row, vocab[, images per row, fixed seq_len])
    "cfg2_clip_l14_336_llama8b": ("clip_l14_336", 4096, 256, 64, 128_258),
    "cfg1_clip_b32_tiny_llama": ("clip_b32", 512, 8, 4, 206),
    # BASELINE configs[2]: SigLIP so400m/14-384 -> Qwen2.5-7B hidden 3584 (64 images/GPU x 8 = 512)
    "cfg3_siglip_so400m_qwen7b": ("siglip_so400m_384", 3584, 64, 64, 152_064),
    # BASELINE configs[3]: interleaved rows, 4 images per 8192-token row (cfg2 encoder), 64 rows = 256 images
    "cfg4_interleaved_8k": ("clip_l14_336", 4096, 256, 0, 128_258, 4, 8192),
    # BASELINE configs[4]: expert-routed image modality (cookbook moe/alignment/avg/shared): 5 CLIP-B/32 experts run
    # on every image, gating-weighted average, shared projector into hidden 4096; RAW mixed-resolution uint8 images
    # ({224,336,512,1024}^2) resized on the device; gating weights are synthetic (the ResNet-50 gate is §8(f) rank 3)
    "cfg5_moe_5x_clip_b32_llama8b": ("clip_b32", 4096, 256, 64, 128_258),
}
MOE_EXPERTS = {"cfg5_moe_5x_clip_b32_llama8b": 5}
# the global batches BASELINE.json:configs names (strong scaling splits these over the ranks)
GLOBAL_BATCH = {"cfg1_clip_b32_tiny_llama": 8, "cfg2_clip_l14_336_llama8b": 256, "cfg3_siglip_so400m_qwen7b": 512,
                "cfg4_interleaved_8k": 256, "cfg5_moe_5x_clip_b32_llama8b": 256}

# dram__bytes_read.sum + dram__bytes_write.sum per launch from the `ncu --set full` capture of this command at
# cfg2 B=256 (profiles/r02_ncu_full_summary.txt, round 2 code): GEMM = mean over one encoder layer's four launches
# (fc2+stats 1.847, qkv(LN folded) 1.184, out-proj+stats 0.883, fc1(LN folded) 1.486 GB); attention3 1.183 GB.
NCU_TRAFFIC = {"cfg2_clip_l14_336_llama8b": {"gemm": 1.350e9, "attention": 1.1833e9, "layernorm": 0.5464e9}}
# the two standalone splice passes, from the round-3 `--set full` capture of tools/splice_roofline.py (same shapes as the
# loops below; profiles/r03_ncu_splice.txt): scatter 1.2088 GB read + 1.1616 GB written for 2.416 GB algorithmic,
# fused embed+splice 4.2237 + 4.2434 GB for 8.590 GB — no re-reads; the part of the last writes that is still in the
# 126 MB L2 when the kernel ends accounts for the 1.5-2 % below algorithmic.
NCU_TRAFFIC_SPLICE = {"scatter_cfg2": 2.3703e9, "embed_splice_cfg4": 8.4671e9}
# spec-sheet peaks (dense bf16 tensor, HBM3e) — SURVEY §8 D3 asks for the fraction of these next to the measured ones
SPEC_TF, SPEC_HBM_GBS = 2250.0, 8000.0


def workload(wl):
    w = WORKLOADS[wl]
    return dict(tower=w[0], H=w[1], images=w[2], n_text=w[3], V=w[4], per_row=w[5] if len(w) > 5 else 1,
                seq_len=w[6] if len(w) > 6 else None)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None
        self.t0 = self.t1 = None   # wall-clock window of the timed region (samples outside it are dropped)

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def wait_first_sample(self, timeout=20.0):
        """nvidia-smi start-up takes seconds and briefly serialises with kernel launches: keep it OUT of the timed
        region by starting the sampler early and waiting for its first line."""
        t = time.time()
        while self.proc is not None and not self.rows and time.time() - t < timeout:
            time.sleep(0.05)

    def mark_begin(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [r for (ts, r) in self.rows if self.t0 is None or (self.t0 <= ts <= (self.t1 or ts) + 0.15)]
        for r in rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------------------------
def build_inputs(wl, seed, images, torch, synth, TOWERS):
    w = workload(wl)
    tower, H, n_text, V, per_row = w["tower"], w["H"], w["n_text"], w["V"], w["per_row"]
    cfg = TOWERS[tower]
    S = w["seq_len"] or cfg.num_patches + 2 + n_text
    assert images % per_row == 0, "images per GPU must be a multiple of the images per row"
    att, start, end, pad = V - 2, V - 4, V - 3, V - 1   # the added special tokens sit at the end of the vocabulary
    ids, _ = synth.make_sequences(images // per_row, S, per_row, cfg.num_patches, V - 4, att, start, end, pad, seed=seed,
                                  min_len=(6144 if w["seq_len"] else None))
    px = synth.make_pixels(cfg, images, seed=seed)
    return cfg, H, V, S, att, ids, px


def cpu_reference_leg(wl, sample_images, steps, warmup, threads, dtype_name="float32", best_of=False, budget_s=None):
    """The reference's CPU path (oracle port: same torch ops as HF CLIP + reference projector + splice) on the host
    cores, on a bounded sample of the workload.  Returns (tokens/s, ms_per_step, description).  `best_of`: the figure
    is the best single step (best of 5 after 2 warm-ups) instead of the mean."""
    import torch
    from multimeditron_b200 import synth
    from multimeditron_b200.configs import TOWERS
    from oracle import splice_oracle, vit_oracle
    torch.set_num_threads(threads)
    cfg, H, V, S, att, ids, px = build_inputs(wl, 0, sample_images, torch, synth, TOWERS)
    dt = getattr(torch, dtype_name)
    sd = {k: v.to(dt) for k, v in synth.make_modality_state_dict(cfg, H, seed=0).items()}
    table = (torch.randn(V, H, generator=torch.Generator().manual_seed(1)) * 0.02).to(dt)
    per_row = workload(wl)["per_row"]
    ranges = [splice_oracle.compute_token_range(ids[b].numpy(), [cfg.num_patches] * per_row, att)
              for b in range(ids.shape[0])]
    bidx, trng = map(torch.from_numpy, splice_oracle.collator_indices(ranges))
    inputs = [px[i] for i in range(sample_images)]

    def step():
        with torch.no_grad():
            proj = vit_oracle.image_modality_forward(sd, cfg, inputs)
            return splice_oracle.embed_modalities_with_text(table, ids, bidx, trng, proj)

    for _ in range(warmup):
        step()
    times = []
    t_begin = time.perf_counter()
    for i in range(steps):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
        if budget_s is not None and i >= 4 and time.perf_counter() - t_begin > budget_s:
            break      # bounded sample: never fewer than 5 timed steps
    steps = len(times)
    dt_s = min(times) if best_of else sum(times) / len(times)
    toks = sample_images * cfg.num_patches / dt_s
    desc = (f"{sample_images} of {WORKLOADS[wl][2]} images per step ({dtype_name}, torch CPU ops, {threads} threads, "
            f"{'best' if best_of else 'mean'} of {steps} after {warmup} warm-ups), full path incl. embedding+splice of those rows")
    return toks, dt_s * 1e3, desc


def cpu_unmodified_reference_leg(wl, sample_images, steps, warmup, threads, dtype_name, budget_s=None):
    """The UNMODIFIED reference on the host cores: the reference's own `MultiModalModelForCausalLM`
    (oracle/_ref staging, or /root/reference in the build container) with its stock `meditron_clip` ImageModality built
    from a random-init HF CLIP checkpoint of the workload's tower and a 1-layer Llama config of the workload's hidden size
    and vocabulary; the timed call is its own `embed_modalities_with_text` (model.py:410-446) on the collator's format.
    Same sample, same best-of reporting as the port leg.  Raises if the reference is not staged."""
    import shutil

    import torch
    from multimeditron_b200 import synth
    from multimeditron_b200.configs import TOWERS
    from oracle import ref_harness, ref_shim, splice_oracle
    ref = ref_shim.import_reference()
    torch.set_num_threads(threads)
    cfg, H, V, S, att, ids, px = build_inputs(wl, 0, sample_images, torch, synth, TOWERS)
    per_row = workload(wl)["per_row"]
    clip_dir = ref_harness.save_clip_checkpoint(cfg, seed=0)
    llama_dir = ref_harness.save_tiny_llama(V, hidden=H, layers=1, heads=max(1, H // 128))
    try:
        mcfg = ref.modalities.AutoModality.config_from_dict(dict(model_type="meditron_clip", hidden_size=H,
                                                                 clip_name=clip_dir))
        rcfg = ref.model.MultimodalConfig(vocab_size=V, modalities=[mcfg], pad_token_idx=V - 1, eos_token_idx=V - 1,
                                          padding_side="right", llm_path=llama_dir, dtype=dtype_name)
        torch.manual_seed(0)
        model = ref.model.MultiModalModelForCausalLM(rcfg).to(getattr(torch, dtype_name)).eval()
    finally:
        shutil.rmtree(clip_dir, ignore_errors=True)
        shutil.rmtree(llama_dir, ignore_errors=True)
    ranges = [splice_oracle.compute_token_range(ids[b].numpy(), [cfg.num_patches] * per_row, att)
              for b in range(ids.shape[0])]
    bidx, trng = map(torch.from_numpy, splice_oracle.collator_indices(ranges))
    pmi = {"stacked": {"image": [px[i] for i in range(sample_images)]}, "batch_idx": {"image": bidx},
           "token_range": {"image": trng}}

    def step():
        with torch.no_grad():
            return model.embed_modalities_with_text(ids, pmi)

    for _ in range(warmup):
        step()
    times, t_begin = [], time.perf_counter()
    for i in range(steps):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
        if budget_s is not None and i >= 4 and time.perf_counter() - t_begin > budget_s:
            break
    dt_s = min(times)
    desc = (f"{sample_images} of {WORKLOADS[wl][2]} images per step ({dtype_name}, the unmodified reference's "
            f"MultiModalModelForCausalLM.embed_modalities_with_text with its stock meditron_clip modality, {threads} threads, "
            f"best of {len(times)} after {warmup} warm-ups)")
    return sample_images * cfg.num_patches / dt_s, dt_s * 1e3, desc


def cpu_baseline_block(wl, sample_images, threads, steps=5, warmup=2, budget_s=45.0, unmodified=False):
    """fp32 AND bf16, 8 images, 2 warm-ups, best of 5.  The headline `value` is the FASTER of the two
    (bf16 on hosts with AMX/AVX512-BF16, fp32 elsewhere) so that the GPU/CPU ratio is not flattered."""
    res = {}
    for dn in ("float32", "bfloat16"):
        toks, ms, desc = cpu_reference_leg(wl, sample_images, max(5, steps), max(2, warmup), threads, dn, best_of=True,
                                           budget_s=budget_s)
        res[dn] = dict(value=toks, ms_per_step=ms, sample=desc)
    best = max(res, key=lambda k: res[k]["value"])
    out = {"value": res[best]["value"], "unit": UNIT, "cores": threads, "kind": "port", "dtype": best,
           "sample": res[best]["sample"], "fp32": res["float32"]["value"], "bf16": res["bfloat16"]["value"],
           "statistic": "best single step (not the mean): conservative for the GPU/CPU ratio, the GPU arm reports the mean",
           "note": "one host (rank 0) only: at N>1 the ratio is N GPUs over one host's cores"}
    if unmodified:
        # `--impl reference`: the unmodified reference itself (staged under oracle/_ref) in the dtype the port found faster;
        # it becomes the headline when it runs, the port's figures stay beside it
        try:
            toks, ms, desc = cpu_unmodified_reference_leg(wl, sample_images, max(5, steps), max(2, warmup), threads, best,
                                                          budget_s=budget_s)
            out.update(port_value=out["value"], port_sample=out["sample"], value=toks, sample=desc, kind="reference")
        except Exception as e:  # noqa: BLE001 — reference not staged / cannot be built on this box: keep the port
            out["reference_unavailable"] = f"{type(e).__name__}: {e}"[:300]
    return out


def torch_gpu_leg(wl, images, steps, warmup, dev, seed=100):
    """The SAME reference path through stock PyTorch on the same B200 ("the real
    number to beat"): HF-CLIP arithmetic as torch dispatches it (cuDNN conv, cuBLASLt GEMMs, SDPA flash/cuDNN attention,
    ATen LayerNorm / elementwise), the reference projector, nn.Embedding and the index_put_ splice — bf16 modules, no
    autocast (regime (i)), inputs resident in HBM.  oracle/vit_oracle.py is bit-identical to the reference modules on
    the same device/dtype (tests/test_oracle_golden.py).  Returns a dict; a second figure times the reference's own
    ingest (list of per-image CPU tensors -> torch.stack -> pageable .to(device), image_modality.py:131-132)."""
    import torch
    from multimeditron_b200 import synth
    from multimeditron_b200.configs import TOWERS
    from oracle import splice_oracle, vit_oracle
    cfg, H, V, S, att, ids, px = build_inputs(wl, seed, images, torch, synth, TOWERS)
    sd = {k: v.to(dev, torch.bfloat16) for k, v in synth.make_modality_state_dict(cfg, H, seed=0).items()}
    table = (torch.randn(V, H, generator=torch.Generator().manual_seed(1)) * 0.02).bfloat16().to(dev)
    per_row = workload(wl)["per_row"]
    ranges = [splice_oracle.compute_token_range(ids[b].numpy(), [cfg.num_patches] * per_row, att)
              for b in range(ids.shape[0])]
    bidx, trng = map(torch.from_numpy, splice_oracle.collator_indices(ranges))
    ids_d, bidx_d, trng_d, px_d = ids.to(dev), bidx.to(dev), trng.to(dev), px.to(dev)
    px_list_h = [px[i] for i in range(images)]          # the collator's format (data_loader.py:144-152)

    @torch.no_grad()
    def step_dev():
        proj = vit_oracle.image_modality_forward(sd, cfg, px_d)
        return splice_oracle.embed_modalities_with_text(table, ids_d, bidx_d, trng_d, proj)

    @torch.no_grad()
    def step_e2e():
        pix = torch.stack(px_list_h, dim=0).to(dev)                       # image_modality.py:131-132
        proj = vit_oracle.image_modality_forward(sd, cfg, pix)
        out = splice_oracle.embed_modalities_with_text(table, ids.to(dev), bidx.to(dev), trng.to(dev), proj)
        return out.view(-1, H)[-1].cpu()

    def timed(fn, n):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    for _ in range(max(2, warmup)):
        step_dev()
    ms = timed(step_dev, steps)
    step_e2e()
    ms_e2e = timed(step_e2e, max(2, steps // 2))
    toks = images * cfg.num_patches
    return {"value": toks / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "images": images,
            "kind": f"torch {torch.__version__} bf16 (cuDNN conv, cuBLASLt, SDPA, index_put_), same path, same B200, "
                    f"inputs resident in HBM",
            "e2e": {"value": toks / (ms_e2e * 1e-3), "ms_per_step": ms_e2e,
                    "note": "the reference's ingest: list of CPU tensors -> torch.stack -> pageable .to(device)"},
            "h2d_bytes_per_step": int(px.numel() * 4 + ids.numel() * 8 + 2 * bidx.numel() * 8)}


def torch_gpu_moe_leg(wl, images, E, steps, warmup, dev, seed=100):
    """cfg5 through stock PyTorch on the same B200: the reference MOEImageModality arithmetic (every expert's HF-CLIP
    tower, ResNet-50 gate as torchvision runs it — unfused BatchNorm —, gate-weighted average, shared projector,
    embedding + index_put_ splice), bf16, pixel_values resident in HBM."""
    import torch
    from multimeditron_b200 import synth
    from multimeditron_b200.configs import TOWERS
    from multimeditron_b200.gating import ResNet50Gate
    from oracle import splice_oracle, vit_oracle
    cfg, H, V, S, att, ids, px = build_inputs(wl, seed, images, torch, synth, TOWERS)
    sd = {k: v.to(dev) for k, v in synth.make_moe_state_dict([cfg] * E, H, False, seed=0, dtype=torch.bfloat16).items()}
    table = (torch.randn(V, H, generator=torch.Generator().manual_seed(1)) * 0.02).bfloat16().to(dev)
    torch.manual_seed(7)
    gate = ResNet50Gate(E, device=dev, fold_bn=False)
    ranges = [splice_oracle.compute_token_range(ids[b].numpy(), [cfg.num_patches], att) for b in range(ids.shape[0])]
    bidx, trng = map(torch.from_numpy, splice_oracle.collator_indices(ranges))
    ids_d, bidx_d, trng_d, px_d = ids.to(dev), bidx.to(dev), trng.to(dev), px.to(dev).bfloat16()
    perm = torch.randperm(E, generator=torch.Generator().manual_seed(5)).tolist()

    @torch.no_grad()
    def step():
        w = gate(px_d)[2]
        proj = vit_oracle.moe_modality_forward(sd, [cfg] * E, px_d, w, perm, "weighted_average", False)
        return splice_oracle.embed_modalities_with_text(table, ids_d, bidx_d, trng_d, proj)

    for _ in range(max(2, warmup)):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"value": images * cfg.num_patches / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "images": images,
            "kind": f"torch {torch.__version__} bf16: {E} HF-CLIP towers + torchvision ResNet-50 gate (cuDNN, unfused BN) + "
                    f"weighted average + projector + index_put_ splice; pixel_values resident in HBM"}


def run_torch_gpu(args):
    """`--impl torch_gpu`: the stock-PyTorch arm alone, one JSON line (rank 0 only under torchrun)."""
    import torch
    if int(os.environ.get("RANK", 0)) != 0:
        return
    assert torch.cuda.is_available()
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    sampler = ClockSampler(dev.index)
    sampler.start()
    w = workload(args.workload)
    images = args.images_per_gpu or w["images"]
    sampler.wait_first_sample()
    sampler.mark_begin()
    r = torch_gpu_leg(args.workload, images, args.steps, max(3, args.warmup), dev)
    sampler.mark_end()
    line = {"impl": "torch_gpu", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": 1, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic (seeded N(0,1) pixels, random-init weights)",
            "config": {"workload": args.workload, "tower": w["tower"], "llm_hidden": w["H"], "images_per_gpu": images,
                       "kind": r["kind"]},
            "clocks": sampler.stop(),
            "e2e": {"value": r["e2e"]["value"], "unit": UNIT, "ms_per_step": r["e2e"]["ms_per_step"],
                    "h2d_bytes_per_step": r["h2d_bytes_per_step"], "d2h_bytes_per_step": w["H"] * 2,
                    "note": r["e2e"]["note"]},
            "gpu_launches": 0}
    emit(line)


def config_block(wl, images, world, px_bytes=None):
    """The `config` object of the JSON line — built from the workload's constants only, so that the B200 arm and the
    `--impl reference` arm print the SAME object (the driver compares them)."""
    from multimeditron_b200.configs import TOWERS
    w = workload(wl)
    cfg = TOWERS[w["tower"]]
    P, H = cfg.num_patches, w["H"]
    S = w["seq_len"] or P + 2 + w["n_text"]
    rows = images // w["per_row"]
    if px_bytes is None:
        px_bytes = images * 3 * cfg.image_size * cfg.image_size * 4
    return {"workload": wl, "tower": w["tower"], "llm_hidden": H, "images_per_gpu": images,
            "global_images": images * world, "rows_per_gpu": rows, "seq_len": S, "tokens_per_image": P,
            "parallelism": f"sample-shard x{world}",
            "l2": "inputs larger than L2 (pixels %.0f MB, inputs_embeds %.0f MB per step)" % (px_bytes / 1e6, rows * S * H * 2 / 1e6),
            "positions": "device prefix sum", "splice": "fused into the last projector GEMM epilogue"}


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = args.cpu_sample
    # fp32 and bf16 on a bounded sample (8 images per step), the driver's --warmup (2..10) and --steps
    # (at least 5, cut short by a 45 s budget per leg), best single step reported; then the unmodified reference itself
    # (oracle/_ref) in the faster dtype — that one is the line's value when it runs, the port is the fallback
    cb = cpu_baseline_block(args.workload, sample, threads, steps=args.steps, warmup=min(args.warmup, 10), budget_s=45.0,
                            unmodified=True)
    toks, ms = cb["value"], 1e3 * sample * TOWERS_P(args.workload) / cb["value"]
    w = workload(args.workload)
    tower, H, ipg, n_text, V = w["tower"], w["H"], w["images"], w["n_text"], w["V"]
    line = {"impl": "reference", "metric": METRIC, "value": toks, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"float32": "f32", "bfloat16": "bf16"}[cb["dtype"]],
            "data": "synthetic (seeded N(0,1) pixels, random-init weights)",
            # the B200 arm's config, verbatim (its GPU-side notes included): this arm is the same workload on the host cores,
            # on a bounded sample of it that cpu_baseline.sample describes
            "config": config_block(args.workload, args.images_per_gpu or ipg, max(1, args.gpus)),
            "cpu_baseline": cb,
            "e2e": {"value": toks, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def TOWERS_P(wl):
    from multimeditron_b200.configs import TOWERS
    return TOWERS[workload(wl)["tower"]].num_patches


# ------------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from multimeditron_b200 import ops, sharding, synth
    from multimeditron_b200.configs import TOWERS
    from multimeditron_b200.modalities import AutoModality
    from multimeditron_b200.model import B200MultiModalEmbedder

    rank, local_rank, world = sharding.env_rank_world()
    assert torch.cuda.is_available(), "bench.py (B200 arm) needs a CUDA device; there is no CPU fallback"
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    w = workload(args.workload)
    tower, H, images, n_text, V = w["tower"], w["H"], w["images"], w["n_text"], w["V"]
    if args.scaling == "strong":   # the workload's GLOBAL batch (BASELINE configs) split over the ranks
        g = GLOBAL_BATCH[args.workload]
        assert g % (world * w["per_row"]) == 0, (g, world)
        images = g // world
    if args.images_per_gpu:
        images = args.images_per_gpu
    cfg, H, V, S, att, ids, px = build_inputs(args.workload, 100 + rank, images, torch, synth, TOWERS)
    P = cfg.num_patches
    real_info = None
    if args.real_images:
        # real photographs through the REAL CPU image processor (what ImageProcessor.process runs,
        # image_modality.py:96) -> fp32 pixel_values; the device pre-processor must reproduce them bit for bit
        from transformers import CLIPImageProcessor
        raws = synth.load_mock_images(images)
        ip = CLIPImageProcessor(size={"shortest_edge": cfg.image_size},
                                crop_size={"height": cfg.image_size, "width": cfg.image_size})
        px = torch.stack([ip(images=r.numpy(), return_tensors="pt")["pixel_values"][0] for r in raws], 0).contiguous()
        from multimeditron_b200.preprocess import DevicePreprocessor
        dp = DevicePreprocessor.from_hf_processor(ip, cfg)(raws, torch.device("cuda", local_rank), torch.float32)
        real_info = {"source": "tests/golden/mock (the reference's mock_dataset photographs, downscaled)",
                     "sizes": sorted({tuple(r.shape[:2]) for r in raws}),
                     "device_preprocess_bit_exact_vs_hf_cpu": bool(torch.equal(dp.cpu(), px))}

    # weights: replicated on every GPU (same seed), random init of the named architecture
    E = MOE_EXPERTS.get(args.workload, 0)
    def timed_load(m, *a, **k):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m.load_reference_state_dict(*a, **k)
        torch.cuda.synchronize()
        return m, time.perf_counter() - t0

    if E:
        names = [f"expert{i}" for i in range(E)]
        order = torch.randperm(E, generator=torch.Generator().manual_seed(5)).tolist()   # shuffled gating class order
        mcfg = AutoModality.config_from_dict({
            "model_type": "moe_meditron_clip_b200", "hidden_size": H, "expert_clip_names": names,
            "expert_towers": {n: tower for n in names}, "image_processor": tower, "fusion_method": "weighted_average",
            "top_k_experts": 1, "gating_class_names": [names[i] for i in order]})
        sd = synth.make_moe_state_dict([cfg] * E, H, False, seed=0, dtype=torch.bfloat16)
        # the reference's gate ON the path (moe/gating.py:73-89: ResNet-50 -> logits -> softmax), random init like the
        # experts; torchvision/cuDNN (library code: SURVEY §8(f) rank 3 keeps it there), bf16 channels-last
        from multimeditron_b200.gating import B200ResNet50Gate, ResNet50Gate
        torch.manual_seed(7)
        tv = ResNet50Gate(E, top_k=1, device="cpu", dtype=torch.float32, fold_bn=False)     # torchvision init = the reference's
        gsd = {"resnet." + k: v for k, v in tv.resnet.state_dict().items()}
        if os.environ.get("MM_B200_GATE", "native") == "cudnn":
            gate = ResNet50Gate(E, top_k=1, device=dev, state_dict=gsd)                      # library path (A/B)
        else:
            gate = B200ResNet50Gate(gsd, E, top_k=1, device=dev)                             # every conv on the tcgen05 GEMM
        del tv
        modality, model_load_s = timed_load(AutoModality.from_config(mcfg), sd, dev, gating=gate)
        # raw mixed-resolution images, packed once (what the collator would hand over), resized on the device
        from multimeditron_b200.preprocess import DevicePreprocessor
        g = torch.Generator().manual_seed(400 + rank)
        sizes = [(224, 336, 512, 1024)[int(i)] for i in torch.randint(0, 4, (images,), generator=g)]
        packed = DevicePreprocessor.for_tower(cfg).pack(
            [torch.randint(0, 256, (n, n, 3), dtype=torch.uint8, generator=g) for n in sizes])
        px = packed.buf            # byte count of the step's image input
    else:
        sd = synth.make_modality_state_dict(cfg, H, seed=0, dtype=torch.bfloat16)
        mcfg = AutoModality.config_from_dict({"model_type": "meditron_clip_b200", "hidden_size": H, "clip_name": tower})
        modality, model_load_s = timed_load(AutoModality.from_config(mcfg), sd, dev)   # pack + LN fold + upload
    del sd
    table = (torch.randn(V, H, generator=torch.Generator().manual_seed(1)) * 0.02).bfloat16().to(dev)
    embedder = B200MultiModalEmbedder(table, [modality], att)

    # ---- device-resident inputs for `value`
    ids_d, px_d = ids.to(dev), (packed.to(dev) if E else px.to(dev))
    pmi_dev = {"stacked": {"image": px_d}, "batch_idx": {}, "token_range": {}}
    # ---- host (pinned) inputs for `e2e`
    ids_h, px_h = ids.pin_memory(), (packed if E else px.pin_memory())
    pmi_host = {"stacked": {"image": px_h}, "batch_idx": {}, "token_range": {}}
    result_h = torch.empty(H, dtype=torch.bfloat16).pin_memory()
    status_h = torch.empty(1, dtype=torch.int32).pin_memory()

    def step_dev():
        # synthetic ids hold exactly one attachment token per modality row (synth.make_sequences), so this leg vouches
        # for the counts and skips the never-observable copy of the attachment token's own embedding row, as the
        # host-id legs do after their eager check
        return embedder(ids_d, pmi_dev, positions_from="device", skip_attachment_rows=True)

    def step_e2e():
        out = embedder(ids_h, pmi_host, positions_from="device")       # H2D of ids + pixels inside
        result_h.copy_(out.view(-1, H)[-1], non_blocking=True)          # D2H read of the step's result (last row)
        status_h.copy_(embedder.status(), non_blocking=True)            # + the data-dependent error word
        return out

    # ---- the collator's REAL format (data_loader.py:126,131,144-152): a Python list of per-image pageable CPU tensors,
    # each its own allocation like the processor's output; the modality stacks them into pinned staging (one host
    # memcpy per image) and uploads on the copy stream (h2d.HostStager)
    pmi_list = None
    if not E:
        pmi_list = {"stacked": {"image": [px[i].clone() for i in range(images)]}, "batch_idx": {}, "token_range": {}}
    ids_pageable = ids.clone()

    def step_e2e_list():
        out = embedder(ids_pageable, pmi_list, positions_from="device")
        result_h.copy_(out.view(-1, H)[-1], non_blocking=True)
        status_h.copy_(embedder.status(), non_blocking=True)
        return out

    # ---- raw uint8 host images for `e2e_uint8`: the same call, resize/crop/normalise on the device
    raw_h = torch.randint(0, 256, (images, cfg.image_size, cfg.image_size, 3), dtype=torch.uint8,
                          generator=torch.Generator().manual_seed(300 + rank)).pin_memory()
    pmi_raw = {"stacked": {"image": raw_h}, "batch_idx": {}, "token_range": {}}

    def step_e2e_u8():
        out = embedder(ids_h, pmi_raw, positions_from="device")        # H2D of ids + raw bytes, pre-processing inside
        result_h.copy_(out.view(-1, H)[-1], non_blocking=True)
        status_h.copy_(embedder.status(), non_blocking=True)
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        return sharding.max_over_ranks(e0.elapsed_time(e1) / steps, dev)

    warm = max(3, args.warmup)
    for _ in range(warm):
        step_dev()
    torch.cuda.synchronize()
    embedder.check()

    if sampler:
        sampler.wait_first_sample()
        sampler.mark_begin()
    # ---- timed region (device-resident inputs), per-op CUDA events recorded on the launching stream
    ops.LAUNCHES = 0
    ms = timed(step_dev, args.steps)                # `value`: no per-op instrumentation in this region
    launches_per_step = ops.LAUNCHES // args.steps
    ops.profile_start()                             # second timed region of the same K steps with one CUDA-event pair per
    if E:                                           # op on the launching stream: per-kernel durations for the rooflines
        os.environ["MM_B200_MOE_STREAMS_SAVED"] = os.environ.get("MM_B200_MOE_STREAMS", "1")
        os.environ["MM_B200_MOE_STREAMS"] = "0"     # (experts serialised here: concurrent streams would overlap the events)
    ms_profiled = timed(step_dev, args.steps)
    prof = ops.profile_stop()
    if E:
        os.environ["MM_B200_MOE_STREAMS"] = os.environ.pop("MM_B200_MOE_STREAMS_SAVED")
    # ---- optional: the same step replayed from ONE CUDA graph (launch-bound workloads: cfg1 is 132 launches for
    # ~0.1 ms of math).  Captured after warm-up, so every one-time attribute / workspace allocation already happened;
    # the graph's private pool keeps every pointer baked into a tensor map stable across replays.
    graph_info = None
    if args.cuda_graph:
        ms_plain = ms
        gstream = torch.cuda.Stream(device=dev)
        gstream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(gstream):
            for _ in range(2):
                step_dev()                              # per-stream workspaces of the capture stream
            gstream.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=gstream):
                g_out = step_dev()
        torch.cuda.current_stream(dev).wait_stream(gstream)
        torch.cuda.synchronize()
        ref_out = step_dev().clone()
        graph.replay()
        torch.cuda.synchronize()
        same = bool(torch.equal(g_out, ref_out))
        ms_graph = timed(graph.replay, max(args.steps, 20))
        graph_info = {"ms_per_step": ms_graph, "value": images * world * P / (ms_graph * 1e-3),
                      "ms_per_step_eager_no_profile": ms_plain, "replay_equals_eager": same,
                      "note": "whole embed_modalities_with_text step captured once, replayed per step"}
    # ---- e2e region (host buffers)
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    ms_e2e_list = None
    if pmi_list is not None:
        for _ in range(2):
            step_e2e_list()
        ms_e2e_list = timed(step_e2e_list, args.steps)
    for _ in range(2):
        step_e2e_u8()
    ops.LAUNCHES = 0
    ms_e2e_u8 = timed(step_e2e_u8, args.steps)
    launches_u8 = ops.LAUNCHES // args.steps
    if sampler:
        sampler.mark_end()
    clocks = sampler.stop() if sampler else None
    embedder.check()
    assert int(status_h.item()) == 0

    # ---- splice standalone (BASELINE metric 'splice HBM GB/s'): in-place scatter of B*P rows x H bf16
    proj = torch.randn(images * P, H, device=dev).bfloat16()
    dest_rows, _, _ = ops.attach_positions(ids_d, att, expected_rows=images * P, want_src_of_pos=False)
    emb = ops.embed_gather(table, ids_d)
    for _ in range(3):
        ops.splice_scatter(emb.view(-1, H), proj, dest_rows)
    ms_sp = timed(lambda: ops.splice_scatter(emb.view(-1, H), proj, dest_rows), max(10, args.steps))
    splice_bytes = 2.0 * images * P * H * 2
    del proj, emb
    # ---- fused embed + splice standalone (SURVEY D5 secondary) at cfg4's shape: inputs_embeds [64, 8192, H] bf16 written in
    # ONE pass from ids, the table and the projected stack (mm_embed_splice_fused); 4 images x 576 tokens per sequence
    es = None
    if rank == 0 and not E:
        Bs4, S4, per4, P4 = 64, 8192, 4, 576
        ids4, _ = synth.make_sequences(Bs4, S4, per4, P4, V - 4, att, V - 4, V - 3, V - 1, seed=77)
        ids4 = ids4.to(dev)
        src4 = torch.randn(Bs4 * per4 * P4, H, device=dev).bfloat16()
        _, sop4, _ = ops.attach_positions(ids4, att, expected_rows=src4.shape[0], want_src_of_pos=True)
        out4 = torch.empty(Bs4, S4, H, dtype=torch.bfloat16, device=dev)
        for _ in range(2):
            ops.embed_splice_fused(table, ids4, sop4, src4, out=out4)
        ms_es = timed(lambda: ops.embed_splice_fused(table, ids4, sop4, src4, out=out4), max(5, args.steps // 2))
        es_bytes = 2.0 * Bs4 * S4 * H * 2
        es = {"kernel": "move_rows_kernel<2> (mm_embed_splice_fused, standalone, cfg4 shape: 64 x 8192 x %d bf16)" % H,
              "bound": "hbm", "achieved": es_bytes / (ms_es * 1e-3) / 1e9, "unit": "GB/s", "bytes": es_bytes, "ms": ms_es,
              "traffic": NCU_TRAFFIC_SPLICE["embed_splice_cfg4"] if H == 4096 else None,
              "traffic_src": "ncu --set full of tools/splice_roofline.py (profiles/r03_ncu_splice.txt)"}
        del ids4, src4, sop4, out4

    gate_ms = None
    if E:
        gpx = torch.randn(images, 3, cfg.image_size, cfg.image_size, device=dev).bfloat16()
        for _ in range(2):
            gate(gpx)
        gate_ms = timed(lambda: gate(gpx), max(5, args.steps))
        del gpx
    total_images = images * world
    pk = peaks()
    flops_img = cfg.flops_per_image(H)
    if E:   # every expert's tower on every image + ONE shared projector
        D = cfg.hidden_size
        proj_fl = 2.0 * P * (D * D + D * H + H * H)
        flops_img = E * (flops_img - proj_fl) + proj_fl
    n_gemm, ms_gemm, fl_gemm = prof.get("mm_gemm_bias_act", (0, 0.0, 0.0))
    gemm_tf = fl_gemm / (ms_gemm * 1e-3) / 1e12 if ms_gemm else 0.0
    path_tf = flops_img * images / (ms * 1e-3) / 1e12
    line = {
        "metric": METRIC, "value": total_images * P / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "bf16", "data": ("real photographs (mock_dataset) through the HF image processor, random-init weights" if args.real_images else
                                  "synthetic (seeded N(0,1) pixels, random-init weights; no network for checkpoints)"),
        "config": config_block(args.workload, images, world, px_bytes=px.numel() * px.element_size()),
        "clocks": clocks,
        "e2e": {"value": total_images * P / ((ms_e2e_list or ms_e2e) * 1e-3), "unit": UNIT,
                "ms_per_step": ms_e2e_list or ms_e2e,
                "h2d_bytes_per_step": int(px.numel() * px.element_size() + ids.numel() * 8),
                "d2h_bytes_per_step": int(H * 2 + 4),
                "input_format": ("the collator's: a Python list of per-image pageable CPU fp32 tensors + pageable int64 ids "
                                 "(data_loader.py:144-152); host stack into pinned staging inside the timed region"
                                 if ms_e2e_list else "one packed pinned host buffer of raw images"),
                "d2h_note": "the consumer of inputs_embeds is the on-device LLM: only the last row + the status word are read back"},
        "e2e_prestacked": {"value": total_images * P / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                           "note": "same call, inputs handed over as ONE pre-stacked pinned host tensor (no host stack)"},
        "ms_per_step_profiled": ms_profiled,
        "model_load_s": model_load_s,
        "cuda_graph": graph_info,
        "real_images": real_info,
        "e2e_uint8": {"value": total_images * P / (ms_e2e_u8 * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e_u8,
                      "h2d_bytes_per_step": int(raw_h.numel() + ids.numel() * 8), "d2h_bytes_per_step": int(H * 2 + 4),
                      "gpu_launches_per_step": launches_u8,
                      "note": "same call fed RAW uint8 [H,W,3] host images: resize/crop/normalise run on the device "
                              "(mm_preprocess_images, bit-exact vs the HF processor) instead of on the CPU"},
        "gpu_launches": launches_per_step * args.steps,
        "gpu_launches_per_step": launches_per_step,
        "roofline": {"kernel": "gemm_bf16_tn_kernel<false,ACT,LN,pair> (tcgen05 cta_group::2 GEMM, %d launches/step)" % (n_gemm // max(1, args.steps)),
                     "bound": "tensor", "achieved": gemm_tf, "peak": pk["tf_sust"], "unit": "TFLOP/s",
                     "frac": gemm_tf / pk["tf_sust"],
                     "traffic": NCU_TRAFFIC.get(args.workload, {}).get("gemm") if images == w["images"] else None,
                     "traffic_src": "ncu --set full, mean bytes per launch over one encoder layer's 4 GEMMs "
                                    "(profiles/r02_ncu_full_summary.txt)",
                     "peak_kind": f"{pk['src']} sustained cuBLAS bf16 (kernel timed inside a long step)",
                     "frac_of_burst": gemm_tf / pk["tf_burst"], "frac_of_spec": gemm_tf / SPEC_TF},
        "path_tensor": {"achieved": path_tf, "unit": "TFLOP/s", "frac_of_sustained": path_tf / pk["tf_sust"],
                        "frac_of_burst": path_tf / pk["tf_burst"], "frac_of_spec": path_tf / SPEC_TF,
                        "gflop_per_image": flops_img / 1e9,
                        "note": "algorithmic FLOPs of the whole path / whole step time (attention, LN, patch, splice included)"},
        "roofline_splice": {"kernel": "move_rows_kernel<0> (mm_splice_scatter, standalone)", "bound": "hbm",
                            "achieved": splice_bytes / (ms_sp * 1e-3) / 1e9, "peak": pk["hbm"], "unit": "GB/s",
                            "frac": splice_bytes / (ms_sp * 1e-3) / 1e9 / pk["hbm"], "bytes": splice_bytes,
                            "ms": ms_sp, "frac_of_spec": splice_bytes / (ms_sp * 1e-3) / 1e9 / SPEC_HBM_GBS,
                            "traffic": (NCU_TRAFFIC_SPLICE["scatter_cfg2"]
                                        if (images, P, H) == (256, 576, 4096) else None),
                            "traffic_src": "ncu --set full of tools/splice_roofline.py, same shape "
                                           "(profiles/r03_ncu_splice.txt)"},
        "op_ms_per_step": {k: v[1] / args.steps for k, v in sorted(prof.items())},
        "roofline_attention": attention_roofline(cfg, images * (E or 1), prof, args.steps, clocks),
    }
    if es is not None:
        es["peak"] = pk["hbm"]
        es["frac"] = es["achieved"] / pk["hbm"]
        es["frac_of_spec"] = es["achieved"] / SPEC_HBM_GBS
        line["roofline_embed_splice"] = es
    if E:
        line["config"]["experts"] = E
        line["config"]["inputs"] = "raw uint8 images, sizes drawn from {224,336,512,1024}^2, resized on the device"
        line["config"]["gating"] = ("reference GatingNetwork on the path (ResNet-50 on every image inside the timed step, overlapping "
                                    "the experts' streams): " + ("torchvision on cuDNN, bf16 channels-last"
                                                                 if os.environ.get("MM_B200_GATE", "native") == "cudnn" else
                                                                 "native — row staging + tcgen05 GEMMs with BatchNorm/ReLU folded") +
                                    "; its FLOPs (~8.2 GFLOP/img) are NOT credited in path_tensor")
        line["config"]["expert_streams"] = os.environ.get("MM_B200_MOE_STREAMS", "1") != "0"
        line["gate_ms_standalone"] = gate_ms
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not E:
        threads = os.cpu_count() or 1
        line["cpu_baseline"] = cpu_baseline_block(args.workload, args.cpu_sample, threads, budget_s=12.0)
    if rank == 0 and world == 1 and not args.no_gpu_baseline:
        del embedder, modality
        torch.cuda.empty_cache()
        if E:
            gb = torch_gpu_moe_leg(args.workload, images, E, max(2, min(args.steps, 5)), 2, dev, seed=100 + rank)
            gb["speedup_value"] = line["value"] / gb["value"]
        else:
            gb = torch_gpu_leg(args.workload, images, max(2, min(args.steps, 5)), 2, dev, seed=100 + rank)
            gb["speedup_value"] = line["value"] / gb["value"]
            gb["speedup_e2e"] = line["e2e"]["value"] / gb["e2e"]["value"]
        line["gpu_baseline"] = gb
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def attention_roofline(cfg, images, prof, steps, clocks):
    """The attention kernel is bound by the MUFU pipe (one ex2 per score; 16 per clock and SM), not by the tensor pipe:
    report algorithmic exponentials/s against 148 SMs x 16/clk x the SM clock sampled under load."""
    n, ms, _ = prof.get("mm_attention_fwd", (0, 0.0, 0.0))
    if not ms:
        return None
    T = cfg.num_tokens
    exps = float(images) * cfg.num_attention_heads * cfg.num_hidden_layers * T * T * steps
    mhz = (clocks or {}).get("sm_mhz") or 1965.0
    peak = 148 * 16 * mhz * 1e6
    ach = exps / (ms * 1e-3)
    return {"kernel": "attention_fwd_kernel", "bound": "mufu", "achieved": ach / 1e12, "peak": peak / 1e12,
            "unit": "Texp/s", "frac": ach / peak, "tflops": 4.0 * exps * cfg.head_dim / (ms * 1e-3) / 1e12,
            "note": "algorithmic T*T exponentials per (image, head, layer); padding to 128-row/column tiles is not credited; "
                    "peak = 148 SMs x 16 ex2/clk x median SM clock under load"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "torch_gpu"])
    ap.add_argument("--workload", default="cfg2_clip_l14_336_llama8b", choices=list(WORKLOADS))
    ap.add_argument("--images-per-gpu", type=int, default=0)
    ap.add_argument("--cpu-sample", type=int, default=8, help="images per CPU-baseline step (BASELINE.md §4: 8)")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip the stock-PyTorch-on-the-same-GPU leg")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: images per GPU fixed; strong: the workload's GLOBAL batch split over the ranks")
    ap.add_argument("--cuda-graph", action="store_true", help="also time the step replayed from one CUDA graph")
    ap.add_argument("--real-images", action="store_true",
                    help="pixels = the reference's mock_dataset photographs (tests/golden/mock) through the HF image "
                         "processor, cycled to the batch; the device pre-processor is checked bit-exact against it")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: anything a library prints there (e.g. NCCL's version banner) goes to stderr
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    elif args.impl == "torch_gpu":
        run_torch_gpu(args)
    else:
        run_b200(args)


_JSON_FD = None


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


if __name__ == "__main__":
    main()
