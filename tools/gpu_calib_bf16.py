"""Calibration: how far do bf16-autocast gradients of the torch ResNet drift from fp32 gradients at random init?
(context for the tolerance of tests/test_gpu_resnet.py)"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.models.nn_compat import build_module
torch.backends.cudnn.allow_tf32 = False
torch.manual_seed(0)
dev = torch.device("cuda")
m = build_module("cifar").to(dev).train()
x = torch.rand(64, 3, 32, 32, device=dev)
y = torch.randint(0, 10, (64,), device=dev)
def grads(autocast):
    m.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        loss = torch.nn.functional.cross_entropy(m(x).float(), y)
    loss.backward()
    return {n: p.grad.clone() for n, p in m.named_parameters()}, loss.item()
g32, l32 = grads(False)
g16, l16 = grads(True)
print("loss", l32, l16)
for n in g32:
    if "conv" in n or "linear" in n or "shortcut.0" in n:
        cos = torch.nn.functional.cosine_similarity(g32[n].flatten(), g16[n].flatten(), dim=0).item()
        print(f"{n:32s} cos {cos:.4f} ratio {(g16[n].norm()/g32[n].norm()).item():.3f}")
