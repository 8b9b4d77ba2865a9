"""MnistNet rounds: native executor (bf16 tcgen05 convolutions) vs the fp32 torch backend on the same GPU, same
configuration as tests/test_gpu_sim.py::test_mnist_rounds_native_backend, a few seeds.  Prints loss / accuracy /
backdoor accuracy per round for both, to calibrate that test's bands."""
import os, sys, yaml, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dba_mod_b200.config import Params
from dba_mod_b200.fl.simulator import Simulator

for seed in (1, 2, 3):
    for backend in ("cuda", "torch"):
        raw = yaml.safe_load(open(os.path.join(ROOT, "utils", "mnist_params.yaml")))
        raw.update(resumed_model=False, synthetic_data=True, synthetic_train_size=6000, synthetic_test_size=1000,
                   number_of_total_participants=20, adversary_list=[3, 7, 11, 15], internal_epochs=4, epochs=5, eta=1.0,
                   lr=0.05, batch_size=32, scale_weights_poison=5, seed=seed, backend=backend,
                   **{"0_poison_epochs": [5], "1_poison_epochs": [5], "2_poison_epochs": [99], "3_poison_epochs": [99]})
        sim = Simulator(Params(raw), device=torch.device("cuda", 0), quiet=True, write_files=False)
        reps = sim.run()
        name = f"{sim.backend.name}/native_train={getattr(sim.backend, 'native_train', False)}"
        sim.close()
        print(f"seed {seed} {name:28s} loss " + " ".join(f"{r.global_loss:6.3f}" for r in reps) + "  acc " +
              " ".join(f"{r.global_acc:5.1f}" for r in reps) + "  asr " + " ".join(f"{r.poison_acc:5.1f}" for r in reps),
              flush=True)
