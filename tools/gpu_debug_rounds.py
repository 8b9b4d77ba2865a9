import os, sys, torch, yaml
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dba_mod_b200.config import Params
from dba_mod_b200.fl.simulator import Simulator
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import bench_params
for graphs in (False, True):
    raw = bench_params(3, 5)
    raw.update(synthetic_train_size=20000, synthetic_test_size=2000, cuda_graphs=graphs, epochs=9)
    sim = Simulator(Params(raw), quiet=True, write_files=False)
    for ep in range(1, 10):
        r = sim.run_round(ep)
        print(f"graphs={graphs} round {ep} adv={r.adversaries} acc {r.global_acc:.2f} asr {r.poison_acc:.2f} steps {r.steps}", flush=True)
    sim.close()
