"""Diagnostics: one native training run vs torch, reporting NaNs / per-layer forward agreement."""
import os, sys
import torch, yaml
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_gpu_resnet import _setup, _one_step_schedule
from dba_mod_b200.fl.native_resnet import NativeResNet
from dba_mod_b200.fl.torch_backend import TorchBackend
from dba_mod_b200.ops.layout import from_planar

torch.backends.cudnn.allow_tf32 = False
p, dev, data, spec, flat = _setup(train=1280, test=200, sampling_dirichlet=False, **{"0_poison_epochs": [1], "internal_epochs": 1,
                                                           "internal_poison_epochs": 1, "lr": 0.0, "poison_lr": 0.0})
names = [17, 3, 5]
sch = _one_step_schedule(p, data, names, dev)
print("T", sch.T, "steps", sch.steps_host, "nvalid", (sch.idx >= 0).sum(2).tolist())
K = len(names)
torch.manual_seed(0)
W0 = flat[None].repeat(K, 1)
W0[:, :spec.Dp] += 0.01 * torch.randn(K, spec.Dp, device=dev)
cnt = torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)
tb = TorchBackend(p, spec, data, dev)
Wt = W0.clone()
rt = tb.train(sch, Wt, cnt.clone(), W0.clone(), True)
nx = NativeResNet(p, spec, data, dev)
Wn = W0.clone()
rn = nx.train(sch, Wn, cnt.clone(), W0.clone(), True)
torch.cuda.synchronize()
print("loss torch", rt.loss_sum.tolist(), "native", rn.loss_sum.tolist())
print("correct", rt.correct.tolist(), rn.correct.tolist(), "count", rt.count.tolist(), rn.count.tolist())
t = nx._train_bufs[K]
B = nx.B
for l in nx.layers:
    r = l.H // l.sub
    y = from_planar(t[("y", l.name)], B, r, r)
    a = from_planar(t[("a", l.name)], B, r, r)
    dy = t[("dy", l.name)].float()
    print(f"{l.name:28s} y nan {int(torch.isnan(y).sum())} absmax {y.abs().max().item():.3e} | a nan {int(torch.isnan(a).sum())} "
          f"absmax {a.abs().max().item():.3e} | dy nan {int(torch.isnan(dy).sum())} absmax {dy.abs().max().item():.3e}")
print("stats nan", int(torch.isnan(t["stats"]).sum()), "saved nan", int(torch.isnan(t["saved"]).sum()),
      "sums nan", int(torch.isnan(t["sums"]).sum()), "Gr nan", int(torch.isnan(t["Gr"]).sum()))
sv = t["saved"].reshape(K, -1, 2)
print("saved mean absmax", sv[..., 0].abs().max().item(), "rstd max", sv[..., 1].max().item(), "rstd min", sv[..., 1].min().item())
for name in spec.param_names:
    e = spec.entries[name]
    a = rt.grad_sum[:, e.offset:e.offset + e.numel]
    b = rn.grad_sum[:, e.offset:e.offset + e.numel]
    cos = torch.nn.functional.cosine_similarity(a, b, dim=1)
    ratio = (b.norm(dim=1) / a.norm(dim=1).clamp(min=1e-12))
    print(f"{name:32s} cos {[round(float(v),4) for v in cos]} ratio {[round(float(v),3) for v in ratio]}")
print("running stats maxdiff", (Wn[:, spec.Dp:] - Wt[:, spec.Dp:]).abs().max().item())
