"""Run one native training step twice on identical inputs and report the first buffer that differs (debug aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, yaml
from dba_mod_b200.config import Params
from dba_mod_b200.data import ImageData
from dba_mod_b200.models import initial_state
from dba_mod_b200.fl.native_resnet import NativeResNet
from dba_mod_b200.fl.schedule import build_schedule, plan_clients

raw = yaml.safe_load(open("utils/cifar_params.yaml"))
raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=1500, synthetic_test_size=200,
           number_of_total_participants=20, cuda_graphs=False, **{"0_poison_epochs": [1], "internal_epochs": 1,
                                                                  "internal_poison_epochs": 3, "lr": 0.02})
p = Params(raw)
dev = torch.device("cuda", 0)
data = ImageData(p, dev)
spec, flat, cnt = initial_state("cifar", 1)
flat = flat.to(dev)
names = [17, 3, 5]
plans = plan_clients(p, data, names, epoch=1)
sch = build_schedule(p, data, plans, 1, dev, torch.Generator(device=dev).manual_seed(5))
K = 3
nx = NativeResNet(p, spec, data, dev)
t = nx._tbufs(K)
t["W"][:, :spec.Df].copy_(flat[None].repeat(K, 1))
t["trig_row"].copy_(sch.trig_row); t["poison_n"].copy_(sch.poison_n)
mask = sch.idx >= 0
t["st_idx"].copy_(sch.idx[:, 0]); t["st_nvalid"].copy_(mask.sum(2).to(torch.int32)[:, 0])
t["st_active"].copy_(sch.active.to(torch.uint8)[:, 0]); t["st_slot"].copy_(sch.epoch_of[:, 0])
t["st_lr"].zero_()
if sch.flip is not None:
    t["st_flip"].copy_(sch.flip.to(torch.uint8)[:, 0])
print("nvalid", t["st_nvalid"].tolist(), "flip", sch.flip is not None)
snaps = []
for it in range(3):
    for key in ("V", "mf", "mi", "nan", "gs"):
        t[key].zero_()
    nx._step_body(K, t, True, sch.flip is not None, False)
    torch.cuda.synchronize()
    snaps.append({k: v.clone() for k, v in t.items() if torch.is_tensor(v)} |
                 {("g", r, i): b.clone() for r in nx.ress for i, b in enumerate(t[("g", r)])})
order = ["xin", "labels"]
for l in nx.layers:
    order += [("y", l.name), ("a", l.name)]
order += ["stats", "saved", "mf", "Gr"]
order += [("dy", l.name) for l in reversed(nx.layers)]
order += ["sums"]
for key in order:
    a, b, c = (s[key].float() for s in snaps)
    d1 = (a - b).abs().max().item(); d2 = (b - c).abs().max().item()
    ref = a.abs().max().item()
    nd = int(((a - b) != 0).sum().item())
    print(f"{str(key):28s} max|a|={ref:10.4g} d(1,2)={d1:10.4g} d(2,3)={d2:10.4g} ndiff={nd}")
