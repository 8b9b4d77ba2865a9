"""Run-to-run noise of the native trainer vs the effect of tail compaction (debug aid)."""
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
           number_of_total_participants=20, **{"0_poison_epochs": [1], "internal_epochs": 1,
                                               "internal_poison_epochs": 3, "lr": float(sys.argv[1]) if len(sys.argv) > 1 else 0.02})
p = Params(raw)
dev = torch.device("cuda", 0)
data = ImageData(p, dev)
spec, flat, cnt = initial_state("cifar", 1)
flat = flat.to(dev)
names = [17, 3, 5]
plans = plan_clients(p, data, names, epoch=1)
sch = build_schedule(p, data, plans, 1, dev, torch.Generator(device=dev).manual_seed(5))
print("steps", sch.steps_host, NativeResNet._segments(list(sch.steps_host)))
K = 3
W0 = flat[None].repeat(K, 1)
outs = {}
for tag, compact, graphs in (("full1", False, True), ("full2", False, True), ("full_nograph", False, False),
                             ("compact", True, True), ("compact_nograph", True, False)):
    p["compact_tail"] = compact
    p["cuda_graphs"] = graphs
    nx = NativeResNet(p, spec, data, dev)
    W, c = W0.clone(), torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)
    r = nx.train(sch, W, c, W0.clone(), True)
    torch.cuda.synchronize()
    outs[tag] = ((W - W0)[:, :spec.Dp], r)
ref = outs["full1"][0]
for tag, (d, r) in outs.items():
    cos = torch.nn.functional.cosine_similarity(ref, d, dim=1)
    print(tag, "cos", [round(x, 5) for x in cos.tolist()], "norm", [round(x, 4) for x in d.norm(dim=1).tolist()],
          "loss", r.loss_sum.flatten().tolist()[:6])
