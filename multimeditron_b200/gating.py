"""The MoE gate of the reference on the device: `GatingNetwork` (src/multimeditron/model/modalities/moe/gating.py:38-89)
= torchvision ResNet-50 with an `fc` of `num_classes` outputs -> logits, top-k indices (computed and discarded by both
callers, image_modality_moe.py:156, image_modality_moe_pep.py:194) and softmax weights.

SURVEY.md §8(f) rank 3 / row A12: the gate is on the path of BASELINE configs[4] but is not one of the kernels
`north_star` names.  Two implementations live here: `B200ResNet50Gate`, the default — the network on this library's
own kernels (csrc/conv_ops.cu row staging + the tcgen05 GEMM with BatchNorm and ReLU folded into its epilogue) — and
`ResNet50Gate`, the same network through torchvision on cuDNN in bf16 channels-last (LIBRARY code; kept as the
comparison partner; bench.py selects it with MM_B200_GATE=cudnn and names the choice in the bench line's `config.gating`).  Either takes the
same `[B,3,224,224]` pixel tensor the experts get and overlaps them on its own stream.
State-dict keys are the reference's: `resnet.*` (inside a MoE modality: `gating_network.resnet.*`)."""
from __future__ import annotations

from typing import Optional, Tuple

import torch


class ResNet50Gate(torch.nn.Module):
    def __init__(self, num_classes: int, top_k: int = 1, device=None, dtype=torch.bfloat16,
                 state_dict: Optional[dict] = None, prefix: str = "resnet.", fold_bn: bool = True):
        super().__init__()
        from torchvision import models                      # gating.py:60-65
        self.resnet = models.resnet50(weights=None)
        self.resnet.fc = torch.nn.Linear(self.resnet.fc.in_features, num_classes)
        self.top_k = int(top_k)
        if state_dict is not None:
            own = {k[len(prefix):]: v for k, v in state_dict.items() if k.startswith(prefix)}
            self.resnet.load_state_dict(own)
        self.resnet.eval()
        if fold_bn:
            # eval-mode BatchNorm is an affine map per channel: fold it into the preceding convolution's weight and bias
            # (in fp32, before the bf16 cast) — one cuDNN conv per layer instead of conv + a memory-bound BN pass
            self.resnet = _fold_batchnorm(self.resnet)
        self.to(device=device, dtype=dtype, memory_format=torch.channels_last)
        self._dtype = dtype

    @torch.no_grad()
    def forward(self, pixel_values: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """gating.py:73-89 -> (logits [B,E], topk_indices [B,k], weights [B,E])."""
        x = pixel_values.to(dtype=self._dtype).contiguous(memory_format=torch.channels_last)
        logits = self.resnet(x)
        _, topk_indices = torch.topk(logits, self.top_k, dim=-1)
        weights = torch.nn.functional.softmax(logits, dim=-1)
        return logits, topk_indices, weights


def _fold_batchnorm(net: torch.nn.Module) -> torch.nn.Module:
    """conv -> bn pairs of a torchvision ResNet (conv1/bn1, every Bottleneck's conv{1,2,3}/bn{1,2,3}, downsample[0]/[1])."""
    from torch.nn.utils.fusion import fuse_conv_bn_eval

    def fold(parent, conv_name, bn_name):
        conv, bn = getattr(parent, conv_name), getattr(parent, bn_name)
        setattr(parent, conv_name, fuse_conv_bn_eval(conv, bn))
        setattr(parent, bn_name, torch.nn.Identity())

    fold(net, "conv1", "bn1")
    for layer in (net.layer1, net.layer2, net.layer3, net.layer4):
        for blk in layer:
            for i in (1, 2, 3):
                fold(blk, f"conv{i}", f"bn{i}")
            if blk.downsample is not None:
                blk.downsample = torch.nn.Sequential(fuse_conv_bn_eval(blk.downsample[0], blk.downsample[1]))
    return net


# ==================================================================================================================
# The same network on the hand-written kernels (no cuDNN): every convolution = row staging + tcgen05 GEMM
# ==================================================================================================================
class _Conv:
    """One convolution with its eval-mode BatchNorm folded in: weight [Cout, Kpad] bf16 (k = (ky*kw+kx)*Cin + c), bias."""

    def __init__(self, w: torch.Tensor, bn: dict, stride: int, pad: int, device, eps: float = 1e-5):
        cout, cin, kh, kw = w.shape
        scale = bn["weight"].double() / torch.sqrt(bn["running_var"].double() + eps)
        wf = (w.double() * scale.view(-1, 1, 1, 1)).permute(0, 2, 3, 1).reshape(cout, kh * kw * cin)     # [Cout, ky,kx,c]
        bf = bn["bias"].double() - bn["running_mean"].double() * scale
        k = wf.shape[1]
        kpad = (k + 63) // 64 * 64
        wp = torch.zeros(cout, kpad, dtype=torch.bfloat16, device=device)
        wp[:, :k] = wf.to(torch.bfloat16).to(device)
        self.w, self.b = wp, bf.to(torch.bfloat16).to(device)
        self.kh, self.kw, self.cin, self.cout, self.stride, self.pad, self.kpad = kh, kw, cin, cout, stride, pad, kpad
        self.pointwise = kh == 1 and kw == 1 and stride == 1 and kpad == cin


class B200ResNet50Gate(torch.nn.Module):
    """`GatingNetwork.forward` (moe/gating.py:73-89) on the sm_100a kernels: torchvision's ResNet-50 (resnet.py, v1.5:
    the stride sits on the 3x3 convolution) with activations kept NHWC = `[B*H*W, C]` bf16, which is the GEMM's A layout.
      * 1x1 stride-1 convolutions: `mm_gemm_bias_act` straight on the activations;
      * 3x3 / strided 1x1 / the 7x7 stem: `mm_conv_rows_*` stages the rows, then the same GEMM;
      * BatchNorm (eval) folded into weight and bias in fp64 at load time; ReLU — and for the last convolution of a
        block `relu(conv + identity)` — in the GEMM epilogue (`MM_ACT_RELU`, applied after the residual add);
      * max-pool / global average pool: `mm_maxpool_nhwc`, `mm_avgpool_nhwc`; the `fc` layer is one more GEMM (its
        `num_classes` rows padded to 8), top-k and softmax over the E logits are two tiny torch calls.
    Takes the same state dict as the reference (`resnet.*`, inside a MoE modality `gating_network.resnet.*`)."""

    LAYERS = ((64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2))       # (planes, blocks, stride) — torchvision resnet50

    def __init__(self, state_dict: dict, num_classes: int, top_k: int = 1, device=None, prefix: str = "resnet."):
        super().__init__()
        dev = torch.device(device if device is not None else "cuda")
        sd = {k[len(prefix):]: v.detach().float().cpu() for k, v in state_dict.items() if k.startswith(prefix)}
        bn = lambda n: {k: sd[f"{n}.{k}"] for k in ("weight", "bias", "running_mean", "running_var")}  # noqa: E731
        self.stem = _Conv(sd["conv1.weight"], bn("bn1"), 2, 3, dev)
        self.blocks = []
        for li, (planes, nblocks, stride) in enumerate(self.LAYERS, start=1):
            for bi in range(nblocks):
                p = f"layer{li}.{bi}."
                s = stride if bi == 0 else 1
                blk = dict(c1=_Conv(sd[p + "conv1.weight"], bn(p + "bn1"), 1, 0, dev),
                           c2=_Conv(sd[p + "conv2.weight"], bn(p + "bn2"), s, 1, dev),
                           c3=_Conv(sd[p + "conv3.weight"], bn(p + "bn3"), 1, 0, dev), down=None)
                if (p + "downsample.0.weight") in sd:
                    blk["down"] = _Conv(sd[p + "downsample.0.weight"], bn(p + "downsample.1"), s, 0, dev)
                self.blocks.append(blk)
        E = int(num_classes)
        assert sd["fc.weight"].shape[0] == E
        n8 = (E + 7) // 8 * 8
        self.fc_w = torch.zeros(n8, sd["fc.weight"].shape[1], dtype=torch.bfloat16, device=dev)
        self.fc_w[:E] = sd["fc.weight"].to(torch.bfloat16).to(dev)
        self.fc_b = torch.zeros(n8, dtype=torch.bfloat16, device=dev)
        self.fc_b[:E] = sd["fc.bias"].to(torch.bfloat16).to(dev)
        self.num_classes, self.top_k, self.device = E, int(top_k), dev

    @staticmethod
    def _conv(c: _Conv, x2d: torch.Tensor, B: int, H: int, W: int, act: int, residual=None):
        """x2d [B*H*W, Cin] -> ([B*Ho*Wo, Cout], Ho, Wo)."""
        from . import ops
        if c.pointwise:
            rows, Ho, Wo = x2d, H, W
        else:
            rows = ops.conv_rows_nhwc(x2d.view(B, H, W, c.cin), c.kh, c.kw, c.stride, c.pad, c.kpad)
            Ho, Wo = ops.conv_out_size(H, c.kh, c.stride, c.pad), ops.conv_out_size(W, c.kw, c.stride, c.pad)
        return ops.gemm_bias_act(rows, c.w, c.b, act=act, residual=residual), Ho, Wo

    @torch.no_grad()
    def forward(self, pixel_values: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        from . import ops
        x = pixel_values.to(self.device)
        if x.dtype not in (torch.float32, torch.bfloat16):
            x = x.float()
        B, _, H, W = x.shape
        c = self.stem
        rows = ops.conv_rows_stem(x.contiguous(), c.kh, c.kw, c.stride, c.pad, c.kpad)
        H, W = ops.conv_out_size(H, c.kh, c.stride, c.pad), ops.conv_out_size(W, c.kw, c.stride, c.pad)
        y = ops.gemm_bias_act(rows, c.w, c.b, act=ops.ACT_RELU)                       # conv1 + bn1 + relu
        y = ops.maxpool_nhwc(y.view(B, H, W, c.cout))                                 # maxpool
        H, W = y.shape[1], y.shape[2]
        y = y.view(B * H * W, c.cout)
        for blk in self.blocks:                                                       # Bottleneck.forward
            identity = y
            o, _, _ = self._conv(blk["c1"], y, B, H, W, ops.ACT_RELU)
            o, Ho, Wo = self._conv(blk["c2"], o, B, H, W, ops.ACT_RELU)
            if blk["down"] is not None:
                identity, _, _ = self._conv(blk["down"], y, B, H, W, ops.ACT_NONE)
            y, _, _ = self._conv(blk["c3"], o, B, Ho, Wo, ops.ACT_RELU, residual=identity)   # relu(bn3(conv3) + identity)
            H, W = Ho, Wo
        feat = ops.avgpool_nhwc(y.view(B, H * W, y.shape[1]))                         # avgpool + flatten
        logits = ops.gemm_bias_act(feat, self.fc_w, self.fc_b)[:, :self.num_classes]  # fc
        _, topk_indices = torch.topk(logits, self.top_k, dim=-1)                      # gating.py:86
        weights = torch.nn.functional.softmax(logits, dim=-1)                         # gating.py:87
        return logits, topk_indices, weights
