"""LOAN tabular data: one client per US state, resident on the device.

Reference: ``LoanHelper.load_data`` walks ``./data/loan/`` and builds one ``StateHelper`` + ``LoanDataset`` per file
``loan_XX.csv`` (state = ``filename[5:7]``), each with its own 80/20 ``train_test_split(random_state=42)`` and its own
shuffled train / ordered test ``DataLoader`` (loan_helper.py:29-78,111-145,148-181).  ``feature_dict`` maps column
name -> column number (from the first file), the benign list is every non-adversary state among the first
``number_of_total_participants`` files, participants = benign + adversaries.

Here all states' rows live in ONE pair of device tensors (``train_x[N, 91]`` fp32 / ``train_y[N]`` and the same for
the test rows); a client is an index range.  ``Mytest`` on LOAN walks every state's test split (test.py:13-27) and
the poison tests stamp every test row of every state without removing the target class (test.py:62-89, Q14), so
``all_test_idx == poison_test_idx`` = all test rows.  Files are read in sorted order (Q24: the reference's order is
whatever ``os.listdir`` returns).  Feature triggers become rows of ``trigger_mask`` / ``trigger_val``
(``triggers.py``).
"""
from __future__ import annotations

import logging
import os
from typing import Dict, List, Tuple

import numpy as np
import torch

from . import synthetic, triggers

logger = logging.getLogger("logger")


def _split(X: np.ndarray, y: np.ndarray) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
    """The reference's per-state split (loan_helper.py:172): sklearn, 20 % test, ``random_state=42``."""
    from sklearn.model_selection import train_test_split
    return train_test_split(X, y, test_size=0.2, random_state=42)


class LoanData:
    def __init__(self, params, device: torch.device) -> None:
        self.params, self.device = params, device
        logger.info("Loading data")
        loan_dir = os.path.join(str(params.get("data_dir", "./data")), "loan")
        files = sorted(f for f in os.listdir(loan_dir) if f.startswith("loan_") and f.endswith(".csv")) \
            if os.path.isdir(loan_dir) else []
        want_syn = params.get("synthetic_data")
        tables: Dict[str, Tuple[np.ndarray, np.ndarray]] = {}
        if want_syn is not True and files:
            import pandas as pd
            columns = None
            for fn in files:
                df = pd.read_csv(os.path.join(loan_dir, fn))
                feats = [c for c in df.columns if c != "loan_status"]
                if columns is None:
                    columns = feats
                tables[fn[5:7]] = (df[feats].to_numpy(dtype=np.float64), df["loan_status"].astype("int64").to_numpy())
            self.is_synthetic = False
        else:
            if want_syn is False:
                raise FileNotFoundError(f"no loan_XX.csv files under {loan_dir!r} and synthetic_data is false")
            sseed = params.get("synthetic_seed")
            tables = synthetic.make_loan(int(params.get("synthetic_train_size") or 60000),
                                         synthetic.DATA_SEED if sseed is None else int(sseed))
            columns = synthetic.loan_feature_names()
            self.is_synthetic = True
        self.state_names: List[str] = list(tables.keys())
        self.feature_names = list(columns)
        self.feature_dict = {n: i for i, n in enumerate(columns)}
        self.n_features = len(columns)
        self.n_classes = synthetic.N_LOAN_CLASSES

        xtr, ytr, xte, yte = [], [], [], []
        self._train_range: Dict[str, Tuple[int, int]] = {}
        self._test_range: Dict[str, Tuple[int, int]] = {}
        a = b = 0
        for st in self.state_names:
            X, y = tables[st]
            x0, x1, y0, y1 = _split(X, y)
            xtr.append(x0); ytr.append(y0); xte.append(x1); yte.append(y1)
            self._train_range[st] = (a, a + len(y0)); a += len(y0)
            self._test_range[st] = (b, b + len(y1)); b += len(y1)
        self.train_x_host = torch.from_numpy(np.concatenate(xtr)).float().contiguous()
        self.train_y_host = torch.from_numpy(np.concatenate(ytr)).long().contiguous()
        self.test_x_host = torch.from_numpy(np.concatenate(xte)).float().contiguous()
        self.test_y_host = torch.from_numpy(np.concatenate(yte)).long().contiguous()
        self.train_x, self.train_y = self.train_x_host.to(device), self.train_y_host.to(device)
        self.test_x, self.test_y = self.test_x_host.to(device), self.test_y_host.to(device)
        self.train_flip = False

        self.all_test_idx = torch.arange(b, dtype=torch.long, device=device)
        self.poison_test_idx = self.all_test_idx                 # Q14: the target class stays in

        # ---- participants (loan_helper.py:131-145)
        self.adversarial_namelist = list(params["adversary_list"])
        self.benign_namelist = [st for st in self.state_names[: int(params["number_of_total_participants"])]
                                if st not in self.adversarial_namelist]
        if params["is_random_namelist"]:
            self.participants_list = self.benign_namelist + self.adversarial_namelist
        else:
            self.participants_list = list(params["participants_namelist"])

        self.trigger_mask, self.trigger_val = (t.to(device) for t in
                                               triggers.all_feature_triggers(params, self.feature_dict, self.n_features))
        self._index: Dict[str, torch.Tensor] = {}

    # ------------------------------------------------------------------------------------------ clients
    def client_size(self, name) -> int:
        lo, hi = self._train_range.get(str(name), (0, 0))
        return hi - lo

    def batches_per_epoch(self, name) -> int:
        B = int(self.params["batch_size"])
        return (self.client_size(name) + B - 1) // B

    def client_index_tensor(self, name) -> torch.Tensor:
        key = str(name)
        t = self._index.get(key)
        if t is None:
            lo, hi = self._train_range.get(key, (0, 0))
            t = torch.arange(lo, hi, dtype=torch.long, device=self.device)
            self._index[key] = t
        return t

    def state_test_index(self, name) -> torch.Tensor:
        lo, hi = self._test_range[str(name)]
        return torch.arange(lo, hi, dtype=torch.long, device=self.device)

    def label_portion(self, name, loan_status: int = 0) -> Tuple[float, float, float]:
        """Share of one label in (all, train, test) rows of a state (``LoanDataset.getPortion``,
        loan_helper.py:197-210)."""
        lo, hi = self._train_range[str(name)]
        tl, th = self._test_range[str(name)]
        tr = int((self.train_y_host[lo:hi] == loan_status).sum())
        te = int((self.test_y_host[tl:th] == loan_status).sum())
        return (tr + te) / max(1, (hi - lo) + (th - tl)), tr / max(1, hi - lo), te / max(1, th - tl)
