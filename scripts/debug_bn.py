import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tutorial-torch-distributed-data-parallel_b200")):
    sys.path.insert(0, p)
import torch, b200ddp
from b200ddp.models import AlexNet, fuse_bn_relu
from oracle import reference_loop as rl
DEV = torch.device("cuda", 0)
for conv in ("cudnn", "b200"):
    torch.manual_seed(1)
    m = AlexNet(bn=True, dropout=0.0, mode="tf32", conv=conv)
    ref = AlexNet(bn=True, dropout=0.0, mode="tf32", conv=conv)
    ref.load_state_dict(m.state_dict())
    m = fuse_bn_relu(b200ddp.SyncBatchNorm.convert_sync_batchnorm(m)).to(DEV)
    ref = ref.to(DEV)
    x, y = rl.synthetic_pool(16, seed=4)
    a = b = x.to(DEV)
    outs_a, outs_b = [], []
    for layer in m.features:
        a = layer(a); outs_a.append((type(layer).__name__, a.clone()))
    for layer in ref.features:
        b = layer(b); outs_b.append((type(layer).__name__, b.clone()))
    for i, ((na, ta), (nb, tb)) in enumerate(zip(outs_a, outs_b)):
        if ta.shape == tb.shape:
            print(conv, i, f"{na:16s} {nb:14s} maxdiff {float((ta-tb).abs().max()):.3e}  scale {float(tb.abs().max()):.3e}")
    la = b200ddp.CrossEntropyLoss()(m(x.to(DEV)), y.to(DEV)); lb = b200ddp.CrossEntropyLoss()(ref(x.to(DEV)), y.to(DEV))
    print(conv, "loss", float(la), float(lb))
