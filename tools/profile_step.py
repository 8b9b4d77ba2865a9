"""Small, ncu-friendly workload: a few training steps of K=10 CIFAR ResNet-18 clients (B=64) and one evaluation
chunk of 10 models x 500 images through the native executor (~600 kernel launches in total).

    python tools/profile_step.py                       # prints CUDA-event timings per phase
    ncu --metrics gpu__time_duration.sum --csv --log-file gpurun_out/launches.csv python tools/profile_step.py
    ncu --set full --import-source on -k regex:conv_tc_kernel -s 60 -c 3 -o gpurun_out/prof python tools/profile_step.py
"""
import os
import sys
import time

import torch
import yaml

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.config import Params  # noqa: E402
from dba_mod_b200.data import ImageData  # noqa: E402
from dba_mod_b200.fl.native_resnet import NativeResNet  # noqa: E402
from dba_mod_b200.fl.schedule import build_schedule, plan_clients  # noqa: E402
from dba_mod_b200.models import initial_state  # noqa: E402

raw = yaml.safe_load(open("utils/cifar_params.yaml"))
raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=6400, synthetic_test_size=1000,
           sampling_dirichlet=False, number_of_total_participants=100, internal_epochs=3, cuda_graphs=False)
p = Params(raw)
dev = torch.device("cuda", 0)
data = ImageData(p, dev)
spec, flat, cnt0 = initial_state("cifar", 1)
flat = flat.to(dev)
names = list(range(10))
plans = plan_clients(p, data, names, epoch=1)
sch = build_schedule(p, data, plans, 1, dev, torch.Generator(device=dev))
K = len(names)
W = flat[None].repeat(K, 1)
cnt = torch.zeros(K, spec.n_counters, dtype=torch.long, device=dev)
nx = NativeResNet(p, spec, data, dev)
nx.eval_chunk = 500


def timed(fn, label, reps=1):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{label}: {e0.elapsed_time(e1) / reps:.3f} ms device, {(time.perf_counter() - t0) * 1e3 / reps:.3f} ms wall",
          flush=True)


print("T", sch.T)
timed(lambda: nx.train(sch, W, cnt, W.clone(), False), f"train {sch.T} steps K={K} (eager)")
timed(lambda: nx.train(sch, W, cnt, W.clone(), False), f"train {sch.T} steps K={K} (eager, 2nd)")
timed(lambda: nx.evaluate(W, "clean", 0), "eval 10 models x 1000 images (2 chunks of 500)")
timed(lambda: nx.evaluate_jobs(W[0], [("clean", 0), ("poison", 0), ("poison", 1), ("poison", 2)]), "eval jobs 1 model x 4 sets")
if os.environ.get("DBA_GRAPH", "1") == "1":
    nx.params["cuda_graphs"] = True
    timed(lambda: nx.train(sch, W, cnt, W.clone(), False), f"train {sch.T} steps K={K} (graph capture round)")
    timed(lambda: nx.train(sch, W, cnt, W.clone(), False), f"train {sch.T} steps K={K} (graph replay)")
print("done")
