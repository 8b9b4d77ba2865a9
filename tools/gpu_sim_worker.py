"""Runs a short CIFAR simulation on the native backend and prints one RESULT json line (rank 0).  Used by the GPU tests
for world-size invariance (launch under torchrun for several ranks).  argv: aggregation method, rounds."""
import json
import os
import sys

import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dba_mod_b200.config import Params  # noqa: E402
from dba_mod_b200.fl.simulator import Simulator  # noqa: E402
from dba_mod_b200.parallel.launcher import make_comm  # noqa: E402

method = sys.argv[1] if len(sys.argv) > 1 else "mean"
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 2
raw = yaml.safe_load(open(os.path.join(ROOT, "utils", "cifar_params.yaml")))
raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=4000, synthetic_test_size=600, epochs=rounds,
           number_of_total_participants=40, no_models=6, aggregation_methods=method, internal_poison_epochs=2,
           internal_epochs=1, adversary_list=[3, 7, 11, 13], cuda_graphs=False)
raw["0_poison_epochs"] = [2]
comm = make_comm()
sim = Simulator(Params(raw), comm=comm, quiet=True, write_files=False)
nccl0 = getattr(sim.comm, "nccl_calls", None)
reps = sim.run()
nccl_in_rounds = (sim.comm.nccl_calls - nccl0) if nccl0 is not None else None
if comm.rank == 0:
    print("RESULT " + json.dumps({"norm": float(sim.G.double().norm()), "acc": reps[-1].global_acc,
                                  "asr": reps[-1].poison_acc, "w": reps[-1].agg_weights,
                                  "path": sim.collective_path, "backend": sim.backend.name,
                                  "nccl_calls_in_rounds": nccl_in_rounds,
                                  "native": [bool(getattr(sim.backend, "native_train", False)),
                                             bool(getattr(sim.backend, "native_eval", False))]}), flush=True)
sim.close()
if comm.world > 1:
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()
