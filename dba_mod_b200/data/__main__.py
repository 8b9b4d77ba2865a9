"""``python -m dba_mod_b200.data [cifar|mnist|tiny|loan]``: the reference's ad-hoc data checks
(image_helper.py:352-377 prints the per-client class histogram of the Dirichlet split, loan_helper.py:212-230 counts
the batches of every state) on a small synthetic set."""
import os
import sys

import torch
import yaml

from ..config import Params
from . import ImageData, LoanData

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main(kind: str = "cifar") -> int:
    raw = yaml.safe_load(open(os.path.join(ROOT, "utils", f"{kind}_params.yaml")))
    raw.update(resumed_model=False, synthetic_data=True)
    if kind == "loan":
        raw.update(synthetic_train_size=12000)
        d = LoanData(Params(raw), torch.device("cpu"))
        total = 0
        for st in d.state_names:
            nb = d.batches_per_epoch(st)
            total += nb
            print(f"{st}: train {d.client_size(st)} rows, {nb} batches, share of label 0: {d.label_portion(st)[0]:.3f}")
        print(f"{len(d.state_names)} states, {total} train batches, {int(d.all_test_idx.numel())} test rows")
        return 0
    raw.update(synthetic_train_size=5000, synthetic_test_size=1000)
    d = ImageData(Params(raw), torch.device("cpu"))
    hist = d.class_histogram()
    for client in list(raw.get("adversary_list", []))[:4] + [0, 1, 2]:
        if client < hist.shape[0]:
            print(f"client {client}: {int(hist[client].sum())} images, per class {hist[client].tolist()}")
    print(f"{hist.shape[0]} clients, sizes min/median/max = {hist.sum(1).min()}/{int(sorted(hist.sum(1))[len(hist) // 2])}/"
          f"{hist.sum(1).max()}; poison test set {int(d.poison_test_idx.numel())} of {int(d.all_test_idx.numel())}")
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1] if len(sys.argv) > 1 else "cifar"))
