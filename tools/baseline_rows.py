"""Print the BASELINE.md section-4 table rows from the bench JSON lines under profiles/r2/."""
import json
import os

R2 = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "r2")


def load(name):
    p = os.path.join(R2, name)
    if not os.path.exists(p):
        return None
    return json.loads(open(p).read().strip().splitlines()[-1])


def cell(d):
    if d is None:
        return "-"
    return f"**{d['value']:.2f}** ({d['ms_per_step']:.0f} ms; e2e {d['e2e']['value']:.2f})"


ref = load("bench_reference_1gpu.json") or load("bench_reference_1gpu_first.json")
rows = [
    ("3. CIFAR-10 ResNet-18, 100/10, DBA, FedAvg (headline)", f"**{ref['value']:.3f}** ({ref['ms_per_step'] / 1e3:.1f} s/round)",
     "bench_own_tf32_1gpu.json", "bench_own_tf32_2gpu.json", "bench_own_tf32_4gpu.json", "bench_own_tf32_8gpu.json",
     "tf32; 8 GPUs: NVLS collectives, global evaluation overlapped"),
    ("3b. same, RFA", "-", "bench_cifar_rfa_1gpu.json", None, None, None, "tf32; attacker weight ~3e-4"),
    ("3c. same, FoolsGold", "-", "bench_cifar_foolsgold_1gpu.json", None, None, None, "tf32"),
    ("1. MNIST MnistNet, 100/10, DBA, FedAvg", "0.108 (9.27 s/round, round-1 measurement, `profiles/bench_other_configs_reference.jsonl`)",
     "bench_mnist_fedavg_1gpu_tf32.json", None, None, None, "tf32 (bf16 mode: 25.4, `bench_mnist_fedavg_1gpu.json`)"),
    ("2. LOAN MLP, 51 clients, DBA feature trigger, FedAvg", "0.121 (8.27 s/round, round-1 measurement)",
     "bench_loan_fedavg_1gpu.json", None, None, None, "fp32 CUDA-core MLP"),
    ("4. CIFAR RFA, 1000 clients / 80 per round", "not run (80 clients per round = ~3 min per reference round)",
     "bench_cifar_rfa_1000_1gpu.json", None, None, "bench_cifar_rfa_1000_8gpu.json", "tf32"),
    ("5. Tiny-ImageNet ResNet-18, FoolsGold, 100/10", "not run (GPU budget)", "bench_tiny_foolsgold_1gpu.json", None, None,
     "bench_tiny_foolsgold_8gpu.json", "tf32, 64x64, 200 classes"),
]
for name, refcell, f1, f2, f4, f8, note in rows:
    print(f"| {name} | {refcell} | " + " | ".join(cell(load(f)) if f else "-" for f in (f1, f2, f4, f8)) + f" | {note} |")
