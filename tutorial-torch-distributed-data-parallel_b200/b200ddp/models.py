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

(* = b200ddp kernels.) Every layer runs b200ddp kernels; `mode` picks their arithmetic for the whole network:
'tf32' (default) = tcgen05 tensor cores with TF32 operands and fp32 accumulation for convolutions and Linear
layers — the reference's own convolution numerics (cuDNN TF32); 'fp32' = strict SIMT fp32 everywhere, the parity
mode of SURVEY 8(d) config 2.

`hidden` scales classifier.1/.4 (BASELINE config 5), `bn=True` builds the AlexNet-BN variant of
SURVEY 8(d) config 4 (BatchNorm2d after each conv, before its ReLU); `SyncBatchNorm.convert_sync_batchnorm`
(README.md:79-81 of the reference) then folds each BN's ReLU into the normalisation kernels.
"""
from __future__ import annotations

import torch
from torch import nn

from . import engine
from . import nn as bnn


class AlexNet(nn.Module):
    def __init__(self, num_classes: int = 10, hidden: int = 4096, dropout: float = 0.5, bn: bool = False,
                 mode="tf32") -> None:
        super().__init__()
        self.bn = bn
        if not bn:
            self.features = nn.Sequential(
                bnn.Conv2d(3, 64, kernel_size=11, stride=4, padding=2, pool=True, mode=mode),
                nn.Identity(), nn.Identity(),
                bnn.Conv2d(64, 192, kernel_size=5, padding=2, pool=True, mode=mode),
                nn.Identity(), nn.Identity(),
                bnn.Conv2d(192, 384, kernel_size=3, padding=1, relu=True, mode=mode),
                nn.Identity(),
                bnn.Conv2d(384, 256, kernel_size=3, padding=1, relu=True, mode=mode),
                nn.Identity(),
                bnn.Conv2d(256, 256, kernel_size=3, padding=1, pool=True, mode=mode),
                nn.Identity(), nn.Identity(),
            )
        else:
            def block(cin, cout, k, s, p, pool):
                # conv -> BatchNorm -> ReLU [-> MaxPool(3, 2)]: ReLU and pooling are b200ddp kernels; after
                # convert_sync_batchnorm the ReLU moves into the SyncBatchNorm kernels (fuse_bn_relu)
                layers = [bnn.Conv2d(cin, cout, kernel_size=k, stride=s, padding=p, mode=mode), nn.BatchNorm2d(cout)]
                layers.append(bnn.ReLUMaxPool2d(relu=True) if pool else bnn.ReLU())
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

    def forward_hidden(self, x: torch.Tensor) -> torch.Tensor:
        """Everything up to (not including) the 10-class head: the [B, hidden] input of classifier.6.
        `b200ddp.evaluate` feeds it to the fused head + metrics kernel."""
        if not torch.is_grad_enabled() and engine.applicable(self, x):
            out = engine.run(self, x, stop_at_hidden=True)      # one native call (csrc/engine.cu)
            if out is not None:
                return out
        return self._hidden_by_modules(x)

    def _hidden_by_modules(self, x: torch.Tensor) -> torch.Tensor:
        convs = [m for m in self.features if isinstance(m, bnn.Conv2d)]
        if convs and x.is_cuda:
            bnn.prepare_conv_weights(convs, first_needs_dgrad=x.requires_grad, in_width=x.shape[3])   # all GEMM weight copies, one launch
        x = self.features(x)
        if x.shape[-2:] != (6, 6):
            x = self.avgpool(x)
        x = torch.flatten(x, 1)
        for layer in list(self.classifier)[:-1]:
            x = layer(x)
        return x

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if engine.applicable(self, x):
            out = engine.run(self, x)           # whole forward (and later backward) as one native call each
            if out is not None:
                return out
        return self.classifier[-1](self._hidden_by_modules(x))


def fuse_bn_relu(model: nn.Module) -> nn.Module:
    """Fold the ReLU that follows a SyncBatchNorm into the normalisation kernels, in every nn.Sequential of `model`:
    `SyncBatchNorm -> ReLU` becomes one fused layer (+ Identity), `SyncBatchNorm -> ReLU+MaxPool` keeps only the
    pooling in the second layer. Called by `SyncBatchNorm.convert_sync_batchnorm`; idempotent."""
    for seq in [m for m in model.modules() if isinstance(m, nn.Sequential)]:
        layers = list(seq.children())
        for i, layer in enumerate(layers[:-1]):
            if not isinstance(layer, bnn.SyncBatchNorm) or layer.fuse_relu:
                continue
            nxt = layers[i + 1]
            if isinstance(nxt, (nn.ReLU, bnn.ReLU)):
                layer.fuse_relu = True
                seq[i + 1] = nn.Identity()
            elif isinstance(nxt, bnn.ReLUMaxPool2d) and nxt.relu:
                layer.fuse_relu = True
                nxt.relu = False
    return model


def load_model(num_classes: int = 10, hidden: int = 4096, seed: int | None = None, mode="tf32", bn: bool = False,
               dropout: float = 0.5) -> AlexNet:
    """Counterpart of reference data_and_toy_model.py:41-45. The reference downloads ImageNet weights
    (no network here): weights are torch's default random init, optionally under `seed`."""
    if seed is not None:
        torch.manual_seed(seed)
    return AlexNet(num_classes=num_classes, hidden=hidden, dropout=dropout, bn=bn, mode=mode)
