"""Calibration of the compute modes against the reference's own precision class.

Per parameter tensor of the CIFAR ResNet-18 at random init (one batch of 64 per client, 3 clients, one of them
poisoning), cosine similarity / norm ratio of the gradient against TRUE fp32 autograd (cuDNN TF32 off) for
  (a) torch with cuDNN TF32 convolutions  = what the reference itself runs (allow_tf32 default)
  (b) this framework, compute_dtype tf32  (fp32 activations, tcgen05 kind::tf32, bf16 hi/lo split weight gradient)
  (c) this framework, compute_dtype bf16  (fast mode)
and, as the noise floor, of two runs of (a) against each other.  Output: profiles/r2/calib_modes.txt.
"""
import os
import sys

import torch
import yaml

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.config import Params  # noqa: E402
from dba_mod_b200.data import ImageData  # noqa: E402
from dba_mod_b200.fl.native_resnet import NativeResNet  # noqa: E402
from dba_mod_b200.fl.schedule import build_schedule, plan_clients  # noqa: E402
from dba_mod_b200.fl.torch_backend import TorchBackend  # noqa: E402
from dba_mod_b200.models import initial_state  # noqa: E402

dev = torch.device("cuda", 0)
torch.backends.cuda.matmul.allow_tf32 = False


def setup(mode):
    raw = yaml.safe_load(open("utils/cifar_params.yaml"))
    raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=1280, synthetic_test_size=200,
               number_of_total_participants=20, sampling_dirichlet=False, compute_dtype=mode, lr=0.0, poison_lr=0.0,
               internal_epochs=1, internal_poison_epochs=1, **{"0_poison_epochs": [1]})
    return Params(raw)


p = setup("tf32")
data = ImageData(p, dev)
spec, flat, _ = initial_state("cifar", 1)
flat = flat.to(dev)
names = [17, 3, 5]
plans = plan_clients(p, data, names, epoch=1)
sch = build_schedule(p, data, plans, 1, dev, torch.Generator(device=dev).manual_seed(5))
K = len(names)
torch.manual_seed(0)
W0 = flat[None].repeat(K, 1)
W0[:, :spec.Dp] += 0.01 * torch.randn(K, spec.Dp, device=dev)
cnt = torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)


def torch_grads(tf32):
    torch.backends.cudnn.allow_tf32 = tf32
    tb = TorchBackend(p, spec, data, dev)
    return tb.train(sch, W0.clone(), cnt.clone(), W0.clone(), True).grad_sum


def native_grads(mode):
    nx = NativeResNet(setup(mode), spec, data, dev)
    return nx.train(sch, W0.clone(), cnt.clone(), W0.clone(), True).grad_sum


ref = torch_grads(False)
cols = {"torch cuDNN-TF32": torch_grads(True), "torch fp32 (2nd run)": torch_grads(False),
        "native tf32": native_grads("tf32"), "native tf32 (2nd run)": native_grads("tf32"),
        "native bf16": native_grads("bf16")}
torch.cuda.synchronize()
lines = ["cosine (min over 3 clients) / worst norm ratio of each parameter gradient vs true-fp32 autograd", ""]
lines.append(f"{'tensor':34s}" + "".join(f"{c:>26s}" for c in cols))
worst = {c: 1.0 for c in cols}
for name in spec.param_names:
    e = spec.entries[name]
    a = ref[:, e.offset:e.offset + e.numel]
    row = f"{name:34s}"
    for c, g in cols.items():
        b = g[:, e.offset:e.offset + e.numel]
        cos = torch.nn.functional.cosine_similarity(a, b, dim=1).min().item()
        ratio = (b.norm(dim=1) / a.norm(dim=1).clamp(min=1e-12))
        dev_ = float((ratio - 1).abs().max())
        worst[c] = min(worst[c], cos)
        row += f"{cos:16.5f} /{dev_:7.4f}"
    lines.append(row)
lines.append("")
lines.append(f"{'worst cosine':34s}" + "".join(f"{worst[c]:26.5f}" for c in cols))
out = "\n".join(lines)
print(out)
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/calib_modes.txt", "w").write(out + "\n")
