"""Measured cost of one lock-step training step vs the number of clients in the batch (calibrates
NativeResNet._STEP_COST).  Usage: python tools/gpu_step_cost.py [cifar|mnist]"""
import os, sys, torch, yaml
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.config import Params
from dba_mod_b200.data import ImageData
from dba_mod_b200.fl.native_resnet import NativeResNet
from dba_mod_b200.fl.schedule import build_schedule, plan_clients
from dba_mod_b200.models import initial_state

KIND = sys.argv[1] if len(sys.argv) > 1 else "cifar"
raw = yaml.safe_load(open(f"utils/{KIND}_params.yaml"))
raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=12800, synthetic_test_size=100,
           sampling_dirichlet=False, number_of_total_participants=20, internal_epochs=2, compact_tail=False,
           overlap_wgrad=os.environ.get("DBA_OVERLAP", "1") == "1")
p = Params(raw)
dev = torch.device("cuda", 0)
data = ImageData(p, dev)
spec, flat, _ = initial_state(KIND, 1)
flat = flat.to(dev)
if KIND == "mnist":
    from dba_mod_b200.fl.native_mnist import NativeMnist
    nx = NativeMnist(p, spec, data, dev)
else:
    nx = NativeResNet(p, spec, data, dev)
for K in (1, 2, 3, 4, 6, 8, 10):
    names = list(range(K))
    plans = plan_clients(p, data, names, epoch=1)
    sch = build_schedule(p, data, plans, 1, dev, torch.Generator(device=dev))
    W = flat[None].repeat(K, 1)
    cnt = torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)
    for _ in range(2):
        nx.train(sch, W.clone(), cnt.clone(), W.clone(), False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    nx.train(sch, W.clone(), cnt.clone(), W.clone(), False)
    e1.record(); torch.cuda.synchronize()
    print(f"K={K:2d}: {sch.T} steps, {e0.elapsed_time(e1) / sch.T:.3f} ms/step", flush=True)
