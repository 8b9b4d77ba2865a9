"""Synthetic data sets "of the named shape" (there is no network on the build / GPU boxes, so torchvision's
``download=True`` of image_helper.py:186-196 and the Lending-Club CSVs of loan_helper.py:117-140 are unavailable).

Images: ``make_images(kind, n_train, n_test, seed)`` returns uint8 NHWC tensors + int64 labels with the shapes of
MNIST (28x28x1, 10 classes), CIFAR-10 (32x32x3, 10) or Tiny-ImageNet (64x64x3, 200).  A sample is its class's smooth
random template, randomly shifted by a few pixels, scaled by a per-sample gain, plus per-pixel noise - learnable by
the reference's networks within a few FL rounds but not trivially separable, and never saturated, so that a pixel
trigger (value 1.0) stays a distinguishable feature.  Both benchmark arms (this framework and the unmodified
reference through ``baseline/ref_harness.py``) draw from this generator with ``DATA_SEED``, i.e. train on identical
bytes.

LOAN: 51 "states" (50 + DC) with 91 numeric feature columns named like the pre-processed Lending-Club table (the six
trigger columns of ``utils/loan_params.yaml`` among them) and a 9-class ``loan_status``.  ``write_loan_csvs`` writes
them in the reference's on-disk layout (``loan_XX.csv``, header row, label column ``loan_status``) so that the
unmodified reference can read them; ``LoanData`` reads the same files or takes the tables in memory.
"""
from __future__ import annotations

import os
from typing import Dict, List, Tuple

import numpy as np
import torch

DATA_SEED = 1234

IMAGE_SHAPES = {           # kind -> (H, W, C, classes, default train size, default test size)
    "mnist": (28, 28, 1, 10, 60000, 10000),
    "cifar": (32, 32, 3, 10, 50000, 10000),
    "tiny-imagenet-200": (64, 64, 3, 200, 100000, 10000),
}
_ALIASES = {"tiny": "tiny-imagenet-200", "tinyimagenet": "tiny-imagenet-200"}


def image_shape(kind: str):
    return IMAGE_SHAPES[_ALIASES.get(kind, kind)]


def _templates(n_classes: int, H: int, W: int, C: int, gen: torch.Generator) -> torch.Tensor:
    """[classes, H, W, C] in [0.2, 0.7]: low-resolution noise, bilinearly enlarged (smooth blobs)."""
    g = max(4, H // 4)
    low = torch.rand(n_classes, C, g, g, generator=gen)
    up = torch.nn.functional.interpolate(low, size=(H + 8, W + 8), mode="bilinear", align_corners=False)
    up = (up - up.amin(dim=(1, 2, 3), keepdim=True)) / (up.amax(dim=(1, 2, 3), keepdim=True)
                                                        - up.amin(dim=(1, 2, 3), keepdim=True) + 1e-6)
    return (0.2 + 0.5 * up).permute(0, 2, 3, 1).contiguous()      # padded by 4 px each side for the shifts


def _draw(templates: torch.Tensor, n: int, H: int, W: int, n_classes: int, noise: float, gen: torch.Generator
          ) -> Tuple[torch.Tensor, torch.Tensor]:
    C = templates.shape[-1]
    y = torch.arange(n) % n_classes
    y = y[torch.randperm(n, generator=gen)]
    x = torch.empty(n, H, W, C, dtype=torch.uint8)
    chunk = 4096
    rows = torch.arange(H)[None, :, None]
    cols = torch.arange(W)[None, None, :]
    for s in range(0, n, chunk):
        yy = y[s:s + chunk]
        m = yy.numel()
        dy = torch.randint(0, 9, (m,), generator=gen)
        dx = torch.randint(0, 9, (m,), generator=gen)
        gain = 0.8 + 0.4 * torch.rand(m, 1, 1, 1, generator=gen)
        t = templates[yy[:, None, None], (rows + dy[:, None, None]), (cols + dx[:, None, None])]   # [m, H, W, C]
        v = t * gain + noise * torch.randn(m, H, W, C, generator=gen)
        x[s:s + m] = (v.clamp_(0.0, 0.9) * 255.0).round_().to(torch.uint8)
    return x, y.long()


def make_images(kind: str, n_train=None, n_test=None, seed: int = DATA_SEED
                ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    """(train_x uint8 [N,H,W,C], train_y int64 [N], test_x, test_y)."""
    H, W, C, n_classes, dtr, dte = image_shape(kind)
    n_train = int(n_train) if n_train else dtr
    n_test = int(n_test) if n_test else dte
    gen = torch.Generator().manual_seed(int(seed))
    tmpl = _templates(n_classes, H, W, C, gen)
    noise = 0.10 if n_classes <= 10 else 0.06
    xtr, ytr = _draw(tmpl, n_train, H, W, n_classes, noise, gen)
    xte, yte = _draw(tmpl, n_test, H, W, n_classes, noise, gen)
    return xtr, ytr, xte, yte


# ------------------------------------------------------------------------------------------------- LOAN
US_STATES = ["AK", "AL", "AR", "AZ", "CA", "CO", "CT", "DC", "DE", "FL", "GA", "HI", "IA", "ID", "IL", "IN", "KS", "KY",
             "LA", "MA", "MD", "ME", "MI", "MN", "MO", "MS", "MT", "NC", "ND", "NE", "NH", "NJ", "NM", "NV", "NY", "OH",
             "OK", "OR", "PA", "RI", "SC", "SD", "TN", "TX", "UT", "VA", "VT", "WA", "WI", "WV", "WY"]

_LOAN_COLUMNS = """loan_amnt funded_amnt funded_amnt_inv pub_rec term int_rate installment grade sub_grade
home_ownership annual_inc verification_status pymnt_plan purpose dti delinq_2yrs inq_last_6mths open_acc revol_bal
revol_util total_acc initial_list_status out_prncp out_prncp_inv total_pymnt total_pymnt_inv total_rec_prncp
total_rec_int total_rec_late_fee recoveries collection_recovery_fee last_pymnt_amnt collections_12_mths_ex_med
policy_code application_type acc_now_delinq tot_coll_amt tot_cur_bal open_acc_6m open_act_il open_il_12m open_il_24m
mths_since_rcnt_il total_bal_il il_util open_rv_12m open_rv_24m max_bal_bc all_util total_rev_hi_lim inq_fi
total_cu_tl inq_last_12m acc_open_past_24mths avg_cur_bal bc_open_to_buy bc_util chargeoff_within_12_mths delinq_amnt
mo_sin_old_il_acct mo_sin_old_rev_tl_op mo_sin_rcnt_rev_tl_op mo_sin_rcnt_tl mort_acc mths_since_recent_bc
mths_since_recent_inq num_accts_ever_120_pd num_actv_bc_tl num_actv_rev_tl num_bc_sats num_bc_tl num_il_tl
num_op_rev_tl num_rev_accts num_rev_tl_bal_gt_0 num_sats num_tl_120dpd_2m num_tl_30dpd num_tl_90g_dpd_24m
num_tl_op_past_12m pct_tl_nvr_dlq percent_bc_gt_75 pub_rec_bankruptcies tax_liens tot_hi_cred_lim total_bal_ex_mort
total_bc_limit total_il_high_credit_limit hardship_flag disbursement_method debt_settlement_flag""".split()

N_LOAN_FEATURES = 91
N_LOAN_CLASSES = 9
# columns that are (almost) always zero in the real table; the shipped triggers set them to 10 ... 100
_RARE = ("num_tl_120dpd_2m", "num_tl_90g_dpd_24m", "pub_rec_bankruptcies", "pub_rec", "acc_now_delinq", "tax_liens",
         "num_tl_30dpd", "chargeoff_within_12_mths", "delinq_amnt", "collections_12_mths_ex_med")


def loan_feature_names() -> List[str]:
    assert len(_LOAN_COLUMNS) == N_LOAN_FEATURES, len(_LOAN_COLUMNS)
    return list(_LOAN_COLUMNS)


def make_loan(n_total: int = 60000, seed: int = DATA_SEED) -> Dict[str, Tuple[np.ndarray, np.ndarray]]:
    """state -> (X float64 [n, 91], y int64 [n]); states hold between ~0.5x and ~4x the average row count."""
    rng = np.random.RandomState(int(seed))
    names = loan_feature_names()
    F = len(names)
    share = rng.dirichlet(np.full(len(US_STATES), 3.0))
    sizes = np.maximum(40, np.round(share * int(n_total)).astype(int))
    # feature scales as after the reference's magnitude rescaling (everything ends up with a mean below ~10)
    scale = rng.uniform(0.3, 4.0, size=F)
    rare = np.array([n in _RARE for n in names])
    Wc = rng.randn(F, N_LOAN_CLASSES) / np.sqrt(F)
    Wc[rare] = 0.0                                    # the label does not depend on the trigger columns
    bias = np.array([2.0, 1.4, 0.2, 0.0, 0.9, -0.2, -0.6, -0.8, -1.0])     # 'Current' / 'Fully Paid' dominate
    out: Dict[str, Tuple[np.ndarray, np.ndarray]] = {}
    for st, n in zip(US_STATES, sizes):
        z = rng.randn(n, F)
        X = np.abs(z) * scale
        X[:, rare] = (rng.rand(n, int(rare.sum())) < 0.03) * rng.randint(1, 3, size=(n, int(rare.sum())))
        logits = 2.5 * (z * (~rare)) @ Wc + bias + 0.3 * rng.randn(1, N_LOAN_CLASSES) + rng.gumbel(size=(n, N_LOAN_CLASSES)) * 0.5
        y = logits.argmax(1).astype(np.int64)
        out[st] = (np.round(X, 4), y)
    return out


def write_loan_csvs(out_dir: str, n_total: int = 60000, seed: int = DATA_SEED) -> List[str]:
    """Per-state CSVs in the reference's layout (``./data/loan/loan_XX.csv``, loan_helper.py:117-124,160-170)."""
    import pandas as pd
    os.makedirs(out_dir, exist_ok=True)
    names = loan_feature_names()
    paths = []
    for st, (X, y) in make_loan(n_total, seed).items():
        df = pd.DataFrame(X, columns=names)
        df["loan_status"] = y
        path = os.path.join(out_dir, f"loan_{st}.csv")
        df.to_csv(path, index=False)
        paths.append(path)
    return paths
