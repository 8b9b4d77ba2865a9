"""dba_mod_b200 — a B200-native federated-learning backdoor-simulation framework.

Capabilities of ehsan886/DBA_mod (many simulated clients per round on MNIST / CIFAR-10 / Tiny-ImageNet / LOAN,
centralized vs distributed (DBA) trigger poisoning with scaled updates, FedAvg / RFA / FoolsGold aggregation,
accuracy + attack-success evaluation, csv_record logging, torch state_dict checkpoints) re-designed for Blackwell:
clients are a batch dimension over flat parameter slabs, data lives on the device, and the hot paths are
hand-written sm_100a kernels (see csrc/ and DESIGN.md).
"""
__version__ = "0.1.0"
