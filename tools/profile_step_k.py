"""ncu-friendly: a few eager training steps at batch width K (argv[1], default 1)."""
import os, sys, torch, yaml
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.config import Params
from dba_mod_b200.data import ImageData
from dba_mod_b200.fl.native_resnet import NativeResNet
from dba_mod_b200.fl.schedule import build_schedule, plan_clients
from dba_mod_b200.models import initial_state

K = int(sys.argv[1]) if len(sys.argv) > 1 else 1
raw = yaml.safe_load(open("utils/cifar_params.yaml"))
raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=1280, synthetic_test_size=100,
           sampling_dirichlet=False, number_of_total_participants=20, internal_epochs=3, compact_tail=False,
           cuda_graphs=False)
p = Params(raw)
dev = torch.device("cuda", 0)
data = ImageData(p, dev)
spec, flat, _ = initial_state("cifar", 1)
flat = flat.to(dev)
nx = NativeResNet(p, spec, data, dev)
names = list(range(K))
plans = plan_clients(p, data, names, epoch=1)
sch = build_schedule(p, data, plans, 1, dev, torch.Generator(device=dev))
W = flat[None].repeat(K, 1)
cnt = torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)
nx.train(sch, W.clone(), cnt.clone(), W.clone(), False)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
nx.train(sch, W.clone(), cnt.clone(), W.clone(), False)
e1.record(); torch.cuda.synchronize()
print(f"K={K}: {sch.T} steps eager, {e0.elapsed_time(e1) / sch.T:.3f} ms/step")
