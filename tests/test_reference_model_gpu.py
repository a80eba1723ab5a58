"""-m gpu + needs_reference: the UNMODIFIED reference model stack driven through the B200 plugin.

What runs is the reference's own code (staged byte for byte under oracle/_ref by oracle/build_ref.py, or
/root/reference in the build container): `MultiModalModelForCausalLM` (model/model.py:205-640), built through its own
`AutoModality` / HF `AutoModel.from_config` route, and `DataCollatorForMultimodal.torch_call`
(model/data_loader.py:31-199) with the reference `ImageProcessor`, a word-level tokenizer carrying the Llama-3 special
tokens, a tiny random-init Llama (d = 512) and a random-init CLIP ViT-B/32 checkpoint — BASELINE configs[0].
The SAME state dict is loaded into two models: `model_type: meditron_clip` (reference forward, HF + cuBLAS/SDPA) and
`model_type: meditron_clip_b200` + `integration.patch_embed_modalities_with_text` (sm_100a kernels).

Tolerances: text rows of inputs_embeds bit-exact; image rows |B200 - fp32 ref| <= 2 x |bf16 ref - fp32 ref| and
<= 2e-2 of range; logits of the (bf16) LLM on top: |B200 - fp32| <= 2 x |bf16 ref - fp32| + 1e-2 of range.
"""
import copy

import pytest
import torch

from multimeditron_b200.configs import TOWERS
from oracle import ref_harness as rh
from oracle import ref_shim

pytestmark = [pytest.mark.gpu, pytest.mark.needs_reference]
DEV = "cuda"
HIDDEN = 512


@pytest.fixture(scope="module")
def stack():
    from multimeditron_b200 import integration
    ref = ref_shim.import_reference()
    cls = integration.register_with_reference()
    tok = rh.make_tokenizer()
    clip = rh.save_clip_checkpoint(TOWERS["clip_b32"])
    llama = rh.save_tiny_llama(len(tok), hidden=HIDDEN, layers=2, heads=8)
    m32 = rh.build_reference_model(ref, clip, llama, tok, "meditron_clip", HIDDEN, "float32", seed=0).to(DEV)
    mbf = copy.deepcopy(m32).to(torch.bfloat16)                                  # the reference's bf16 regime (i)
    mb2 = rh.build_reference_model(ref, clip, llama, tok, "meditron_clip_b200", HIDDEN, "float32", seed=1)
    missing, unexpected = mb2.load_state_dict(m32.state_dict(), strict=True)     # same keys: it is a drop-in
    assert not missing and not unexpected
    mb2 = mb2.to(torch.bfloat16).to(DEV)
    assert isinstance(mb2.modalities_with_projection[0], cls)
    integration.patch_embed_modalities_with_text(mb2, tokenizer=tok, attachment_token=rh.ATTACHMENT_TOKEN)
    imgs = rh.mock_images() + rh.synthetic_images(6, size=200)
    coll = rh.make_collator(ref, tok, m32)
    batch = coll(rh.make_samples(imgs, [1, 0, 2, 1, 3, 1], seed=3))              # 8 images, one row without any
    return dict(ref=ref, cls=cls, tok=tok, m32=m32, mbf=mbf, mb2=mb2, batch=batch, coll=coll, imgs=imgs,
                clip=clip, llama=llama, integration=integration)


def _att_mask(stack):
    att = stack["tok"].convert_tokens_to_ids(rh.ATTACHMENT_TOKEN)
    return (stack["batch"]["input_ids"] == att).to(DEV)


@torch.no_grad()
def test_collator_batch_is_cfg1_shaped(stack):
    b = stack["batch"]
    pmi = b["processed_multimodal_inputs"]
    assert len(pmi["stacked"]["image"]) == 8 and pmi["stacked"]["image"][0].shape == (3, 224, 224)
    assert pmi["batch_idx"]["image"].numel() == 8 * 49 == int(_att_mask(stack).sum())


@torch.no_grad()
def test_embed_modalities_with_text_through_the_plugin(stack):
    b, m32, mbf, mb2 = stack["batch"], stack["m32"], stack["mbf"], stack["mb2"]
    ids, pmi = b["input_ids"], b["processed_multimodal_inputs"]
    e32 = m32.embed_modalities_with_text(ids.to(DEV), pmi).float()
    ebf = mbf.embed_modalities_with_text(ids.to(DEV), pmi).float()
    n_into = stack["cls"].b200_forward_into_calls
    eb2 = mb2.embed_modalities_with_text(ids, pmi)                    # host ids + the collator's list of CPU tensors
    mb2.b200_check()
    assert stack["cls"].b200_forward_into_calls == n_into + 1        # the fused device path really ran
    assert mb2.b200_state["fused_calls"] >= 1 and mb2.b200_state["device_position_calls"] >= 1
    assert eb2.dtype == torch.bfloat16 and eb2.shape == e32.shape
    att = _att_mask(stack)
    assert torch.equal(eb2[~att].float(), ebf[~att])                 # text rows: the embedding rows, bit-exact
    scale = e32[att].abs().max().item()
    err, ref_err = (eb2.float() - e32)[att].abs().max().item(), (ebf - e32)[att].abs().max().item()
    mean_rel = ((eb2.float() - e32)[att].abs().mean() / e32[att].abs().mean()).item()
    print(f"[ref model, cfg1] image rows max-abs {err:.3e} (reference bf16: {ref_err:.3e}) /absmax {err / scale:.3e} "
          f"mean-rel {mean_rel:.3e}")
    assert err <= 2.0 * ref_err and err / scale <= 2e-2
    # the collator's own index vectors instead of the device prefix sum: same bits
    mc = stack["integration"].patch_embed_modalities_with_text(copy.copy(mb2), positions_from="collator")
    ec = mc.embed_modalities_with_text(ids, pmi)
    mc.b200_check()
    assert torch.equal(ec, eb2)
    # regime (ii) of SURVEY C5: the reference in bf16 under torch.autocast (inference.py:88) — LN/softmax in fp32
    with torch.autocast("cuda", dtype=torch.bfloat16):
        eac = mbf.embed_modalities_with_text(ids.to(DEV), pmi).float()
    err_ac = (eb2.float() - eac)[att].abs().max().item()
    print(f"[ref model, cfg1] vs the autocast regime: max-abs {err_ac:.3e} /absmax {err_ac / scale:.3e}")
    assert err_ac / scale <= 2e-2


@torch.no_grad()
def test_forward_logits_and_loss(stack):
    b, m32, mbf, mb2 = stack["batch"], stack["m32"], stack["mbf"], stack["mb2"]
    kw = dict(input_ids=b["input_ids"].to(DEV), attention_mask=b["attention_mask"].to(DEV),
              position_ids=b["position_ids"].to(DEV), labels=b["labels"].to(DEV),
              processed_multimodal_inputs=b["processed_multimodal_inputs"])
    o32, obf = m32(**kw), mbf(**kw)
    ob2 = mb2(**kw)                                                   # model.py:449-526, unchanged signature
    mb2.b200_check()
    scale = o32.logits.abs().max().item()
    err = (ob2.logits.float() - o32.logits).abs().max().item()
    ref_err = (obf.logits.float() - o32.logits).abs().max().item()
    print(f"[ref model, cfg1] logits max-abs {err:.3e} (reference bf16: {ref_err:.3e}); loss {ob2.loss.item():.4f} "
          f"vs fp32 {o32.loss.item():.4f} / bf16 {obf.loss.item():.4f}")
    assert err <= 2.0 * ref_err + 1e-2 * scale
    assert abs(ob2.loss.item() - o32.loss.item()) <= 2.0 * abs(obf.loss.item() - o32.loss.item()) + 2e-2


def test_generate_runs_through_the_plugin(stack):
    """model.py:528-640: `generate` calls embed_modalities_with_text OUTSIDE its no_grad block on a model whose
    parameters all require grad (fresh from_pretrained) — the patched model must still run."""
    ref, tok, mb2 = stack["ref"], stack["tok"], stack["mb2"]
    assert any(p.requires_grad for p in mb2.parameters())
    coll = rh.make_collator(ref, tok, stack["m32"], add_generation_prompt=True)
    batch = coll(rh.make_samples(stack["imgs"], [1, 2], seed=9))
    out = mb2.generate(batch, max_new_tokens=4, do_sample=False)
    mb2.b200_check()
    assert out.shape[0] == 2 and 1 <= out.shape[1] <= 4 and out.dtype == torch.int64
    with torch.no_grad():
        want = stack["mbf"].generate(batch, max_new_tokens=1, do_sample=False)
    print("[ref model] generate:", out.tolist(), "reference bf16 first token:", want.tolist())


def test_device_side_generate_is_token_identical_to_the_reference_loop(stack):
    """integration.b200_generate: the reference's decode loop (model.py:581-638) without the per-token `.cpu()` / Python
    sampling loop.  Greedy decoding must reproduce `generate` token for token on the SAME model, also when the host only
    checks the stop condition every few steps, and rows that hit EOS early must be padded with EOS exactly like :620-622."""
    ref, tok, mb2 = stack["ref"], stack["tok"], stack["mb2"]
    coll = rh.make_collator(ref, tok, stack["m32"], add_generation_prompt=True)
    batch = coll(rh.make_samples(stack["imgs"], [1, 2, 0], seed=11))
    want = mb2.generate(batch, max_new_tokens=12, do_sample=False)
    for sync_every in (1, 5, 16):
        got = mb2.b200_generate(batch, max_new_tokens=12, do_sample=False, sync_every=sync_every)
        assert got.dtype == torch.int64 and not got.is_cuda
        assert torch.equal(got, want), (sync_every, got.tolist(), want.tolist())
    # force an early EOS: make the first generated token of row 0 the EOS id and check the tail handling
    old = mb2.config.eos_token_idx
    try:
        mb2.config.eos_token_idx = int(want[0, 0])
        w2 = mb2.generate(batch, max_new_tokens=8, do_sample=False)
        g2 = mb2.b200_generate(batch, max_new_tokens=8, do_sample=False, sync_every=3)
        assert torch.equal(g2, w2), (g2.tolist(), w2.tolist())
    finally:
        mb2.config.eos_token_idx = old
    s = mb2.b200_generate(batch, max_new_tokens=4, do_sample=True, temperature=1.0)
    assert s.shape == (3, 4) or s.shape[1] <= 4
    mb2.b200_check()


def test_errors_are_the_references_and_do_not_stick(stack):
    mb2, b = stack["mb2"], stack["batch"]
    att = stack["tok"].convert_tokens_to_ids(rh.ATTACHMENT_TOKEN)
    ids = b["input_ids"].clone()
    ids[0, (ids[0] == att).nonzero()[0]] = 1                          # one attachment token fewer than embeddings
    with torch.no_grad():
        with pytest.raises(RuntimeError):                             # host ids: caught before any launch
            mb2.embed_modalities_with_text(ids, b["processed_multimodal_inputs"])
        mb2.embed_modalities_with_text(ids.to(DEV), b["processed_multimodal_inputs"])   # device ids: lazy
        with pytest.raises(RuntimeError):
            mb2.b200_check()
        mb2.b200_check()                                              # cleared by the failing check: no sticky error
        bad = b["input_ids"].clone()
        bad[0, 0] = 10 ** 6
        with pytest.raises(IndexError):
            mb2.embed_modalities_with_text(bad, b["processed_multimodal_inputs"])
    with pytest.raises(ValueError):                                   # device positions need the token id
        stack["integration"].patch_embed_modalities_with_text(copy.copy(mb2))
    # a trainable embedding table with autograd on would silently lose its gradient -> loud error instead
    for p in mb2.parameters():
        p.requires_grad = False
    mb2.get_input_embeddings().weight.requires_grad = True
    with pytest.raises(NotImplementedError):
        mb2.embed_modalities_with_text(b["input_ids"], b["processed_multimodal_inputs"])
    for p in mb2.parameters():
        p.requires_grad = True


def test_alignment_training_step_matches_reference_gradients(stack):
    """freeze_for_alignment (model.py:310-326): only the projector trains.  One forward + backward of the reference
    model through the plugin (ViT on the kernels, projector fwd+bwd = autograd_ops.ProjectorFunction, differentiable
    splice) against the fp32 reference's autograd: loss and the six projector gradients."""
    b, m32, mbf, mb2 = stack["batch"], stack["m32"], stack["mbf"], stack["mb2"]
    kw = lambda: dict(input_ids=b["input_ids"].to(DEV), attention_mask=b["attention_mask"].to(DEV),   # noqa: E731
                      position_ids=b["position_ids"].to(DEV), labels=b["labels"].to(DEV),
                      processed_multimodal_inputs=b["processed_multimodal_inputs"])
    grads = {}
    for tag, m in (("fp32", m32), ("bf16", mbf), ("b200", mb2)):
        m.freeze_for_alignment()
        m.zero_grad(set_to_none=True)
        out = m(**kw())
        out.loss.backward()
        proj = m.modalities_with_projection[0].projector
        grads[tag] = {n: p.grad.detach().float().clone() for n, p in proj.named_parameters()}
        assert all(p.grad is None for p in m.modalities_with_projection[0].feature_extractor.parameters())
        grads[tag]["loss"] = out.loss.detach().float()
    mb2.b200_check()
    for n in grads["fp32"]:
        if n == "loss":
            continue
        g32, gbf, gb2 = grads["fp32"][n], grads["bf16"][n], grads["b200"][n]
        scale = g32.abs().max().item()
        err, ref_err = (gb2 - g32).abs().max().item(), (gbf - g32).abs().max().item()
        print(f"[alignment grads] {n}: max-abs {err:.3e} (reference bf16: {ref_err:.3e}) /absmax {err / scale:.3e}")
        assert err <= 2.0 * ref_err + 2e-2 * scale, n
    for m in (m32, mbf, mb2):
        for p in m.parameters():
            p.requires_grad = True
