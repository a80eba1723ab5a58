"""The MoE gate of the reference on the device: `GatingNetwork` (src/multimeditron/model/modalities/moe/gating.py:38-89)
= torchvision ResNet-50 with an `fc` of `num_classes` outputs -> logits, top-k indices (computed and discarded by both
callers, image_modality_moe.py:156, image_modality_moe_pep.py:194) and softmax weights.

SURVEY.md §8(f) rank 3 / row A12: the gate is on the path of BASELINE configs[4] but is not one of the kernels
`north_star` names; it runs through torchvision on cuDNN here (LIBRARY code, stated as such in DESIGN.md and in the
bench line's `config.gating`), in bf16 channels-last.  It takes the same `[B,3,224,224]` pixel tensor the experts get and
is launched on the caller's stream while the experts run on their side streams, so it overlaps them.
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
