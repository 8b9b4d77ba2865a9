"""ncu-friendly evaluation workload: M CIFAR ResNet-18 models x N test images through the native executor.

    python tools/profile_eval.py [M] [N]
    ncu --metrics gpu__time_duration.sum --csv --log-file gpurun_out/launches_eval.csv python tools/profile_eval.py
"""
import os
import sys

import torch
import yaml

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.config import Params  # noqa: E402
from dba_mod_b200.data import ImageData  # noqa: E402
from dba_mod_b200.fl.native_resnet import NativeResNet  # noqa: E402
from dba_mod_b200.models import initial_state  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 10
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
raw = yaml.safe_load(open("utils/cifar_params.yaml"))
raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=640, synthetic_test_size=N,
           native_eval_chunk=int(os.environ.get("DBA_EVAL_CHUNK", "500")))
p = Params(raw)
dev = torch.device("cuda", 0)
data = ImageData(p, dev)
spec, flat, _ = initial_state("cifar", 1)
W = flat.to(dev)[None].repeat(M, 1)
nx = NativeResNet(p, spec, data, dev)


def timed(fn, label):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{label}: {e0.elapsed_time(e1):.3f} ms", flush=True)


timed(lambda: nx.evaluate(W, "clean", 0), f"eval {M} models x {N} images, chunk {nx.eval_chunk}")
timed(lambda: nx.evaluate_jobs(W[0], [("clean", 0), ("poison", 0), ("poison", 1), ("poison", 2), ("poison", 3),
                                      ("poison", 4)]), f"eval jobs 1 model x 6 sets of {N}")
print("done")
