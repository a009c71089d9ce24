"""M0 — the toy model: torchvision AlexNet with a 10-class head (reference data_and_toy_model.py:41-45).

The module tree reproduces torchvision/models/alexnet.py:17-52 index for index, so the 16 parameter
tensors carry the torchvision names (`features.0.weight` ... `classifier.6.bias`) and a state dict
moves freely between this class and `torchvision.models.alexnet`. Fused kernels take the place of
several stock layers; the layers they absorb stay in the Sequential as `nn.Identity` so indices
(and therefore names) do not shift:

  features:   0 Conv 11x11/4  1 ReLU+MaxPool*  2 (fused)   3 Conv 5x5   4 ReLU+MaxPool*  5 (fused)
              6 Conv 3x3      7 ReLU*          8 Conv 3x3  9 ReLU*     10 Conv 3x3      11 ReLU+MaxPool*  12 (fused)
  avgpool:    AdaptiveAvgPool2d((6,6)) — the identity at 224x224 input, skipped when the map is already 6x6
  classifier: 0 Dropout*      1 Linear+ReLU+Dropout*  2,3 (fused)  4 Linear+ReLU*  5 (fused)  6 Linear* (head)

(* = b200ddp kernels.) The five convolutions are b200ddp implicit-GEMM tcgen05 kernels too (`conv="b200"`,
default; conv 6 and 8 fold their ReLU into the epilogue); `conv="cudnn"` keeps torch.nn.Conv2d for A/B runs.

`hidden` scales classifier.1/.4 (BASELINE config 5), `bn=True` builds the AlexNet-BN variant of
SURVEY 8(d) config 4 (BatchNorm2d after each conv, before its ReLU).
"""
from __future__ import annotations

import torch
from torch import nn

from . import nn as bnn


class AlexNet(nn.Module):
    def __init__(self, num_classes: int = 10, hidden: int = 4096, dropout: float = 0.5, bn: bool = False,
                 mode="tf32", conv: str = "b200") -> None:
        super().__init__()
        self.bn = bn
        if conv not in ("b200", "cudnn"):
            raise ValueError("conv must be 'b200' (own implicit-GEMM kernels) or 'cudnn' (torch.nn.Conv2d)")
        if not bn and conv == "b200":
            self.features = nn.Sequential(
                bnn.Conv2d(3, 64, kernel_size=11, stride=4, padding=2, pool=True),
                nn.Identity(), nn.Identity(),
                bnn.Conv2d(64, 192, kernel_size=5, padding=2, pool=True),
                nn.Identity(), nn.Identity(),
                bnn.Conv2d(192, 384, kernel_size=3, padding=1, relu=True),
                nn.Identity(),
                bnn.Conv2d(384, 256, kernel_size=3, padding=1, relu=True),
                nn.Identity(),
                bnn.Conv2d(256, 256, kernel_size=3, padding=1, pool=True),
                nn.Identity(), nn.Identity(),
            )
        elif not bn:
            self.features = nn.Sequential(
                nn.Conv2d(3, 64, kernel_size=11, stride=4, padding=2),
                bnn.ReLUMaxPool2d(), nn.Identity(),
                nn.Conv2d(64, 192, kernel_size=5, padding=2),
                bnn.ReLUMaxPool2d(), nn.Identity(),
                nn.Conv2d(192, 384, kernel_size=3, padding=1),
                bnn.ReLU(),
                nn.Conv2d(384, 256, kernel_size=3, padding=1),
                bnn.ReLU(),
                nn.Conv2d(256, 256, kernel_size=3, padding=1),
                bnn.ReLUMaxPool2d(), nn.Identity(),
            )
        else:
            def block(cin, cout, k, s, p, pool):
                conv_cls = bnn.Conv2d if conv == "b200" else nn.Conv2d
                layers = [conv_cls(cin, cout, kernel_size=k, stride=s, padding=p), nn.BatchNorm2d(cout), nn.ReLU(inplace=True)]
                if pool:
                    layers.append(nn.MaxPool2d(kernel_size=3, stride=2))
                return layers
            self.features = nn.Sequential(*block(3, 64, 11, 4, 2, True), *block(64, 192, 5, 1, 2, True),
                                          *block(192, 384, 3, 1, 1, False), *block(384, 256, 3, 1, 1, False),
                                          *block(256, 256, 3, 1, 1, True))
        self.avgpool = nn.AdaptiveAvgPool2d((6, 6))
        self.classifier = nn.Sequential(
            bnn.Dropout(p=dropout),
            bnn.Linear(256 * 6 * 6, hidden, relu=True, dropout=dropout, mode=mode), nn.Identity(), nn.Identity(),
            bnn.Linear(hidden, hidden, relu=True, mode=mode), nn.Identity(),
            bnn.Linear(hidden, num_classes, mode=mode),
        )

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        convs = [m for m in self.features if isinstance(m, bnn.Conv2d)]
        if convs and x.is_cuda:
            bnn.prepare_conv_weights(convs, first_needs_dgrad=x.requires_grad)     # all GEMM weight copies, one launch
        x = self.features(x)
        if x.shape[-2:] != (6, 6):
            x = self.avgpool(x)
        x = torch.flatten(x, 1)
        return self.classifier(x)


def fuse_bn_relu(model: AlexNet) -> AlexNet:
    """After convert_sync_batchnorm: fold each `SyncBatchNorm -> ReLU` pair of AlexNet-BN into one kernel."""
    feats = list(model.features.children())
    for i, layer in enumerate(feats[:-1]):
        if isinstance(layer, bnn.SyncBatchNorm) and isinstance(feats[i + 1], nn.ReLU):
            layer.fuse_relu = True
            model.features[i + 1] = nn.Identity()
    return model


def load_model(num_classes: int = 10, hidden: int = 4096, seed: int | None = None, mode="tf32", bn: bool = False,
               dropout: float = 0.5, conv: str = "b200") -> AlexNet:
    """Counterpart of reference data_and_toy_model.py:41-45. The reference downloads ImageNet weights
    (no network here): weights are torch's default random init, optionally under `seed`."""
    if seed is not None:
        torch.manual_seed(seed)
    return AlexNet(num_classes=num_classes, hidden=hidden, dropout=dropout, bn=bn, mode=mode, conv=conv)
