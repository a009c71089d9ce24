import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tutorial-torch-distributed-data-parallel_b200")]
import torch
from b200ddp import nn as bnn
DEV = torch.device("cuda", 0)
torch.manual_seed(0)
x = torch.randn(2, 3, 224, 224, device=DEV)
conv = bnn.Conv2d(3, 64, 11, stride=4, padding=2).to(DEV)
y = conv(x)
torch.cuda.synchronize()
ref = torch.nn.functional.conv2d(x.double(), conv.weight.double(), conv.bias.double(), stride=4, padding=2)
print("err", float((y.double() - ref).abs().max() / ref.abs().max()))
