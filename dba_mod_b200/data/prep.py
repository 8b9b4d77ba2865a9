"""Offline data preparation (run once, before training), same outputs as the reference's scripts:

    python -m dba_mod_b200.data.prep loan  <lending-club loan.csv> [out_dir=./data/loan]
    python -m dba_mod_b200.data.prep tiny  <.../tiny-imagenet-200/val>
    python -m dba_mod_b200.data.prep synthetic-loan [out_dir=./data/loan] [rows=60000]

``preprocess_loan`` (reference utils/loan_preprocess.py:1-58): drop the identifier / free-text / date / hardship /
settlement / joint-application columns, ``fillna(0)``, label-encode every remaining object column in order of first
appearance (``addr_state`` excepted), divide numeric columns by 10 / 100 / 10000 when their mean lies in (10, 100] /
(100, 1000] / above 1000, then split by ``addr_state`` into ``loan_<STATE>.csv`` without that column.  The reference
does this with per-value ``df.loc`` assignments and a Python row loop for writing; here it is vectorised pandas.

``reformat_tiny_val`` (reference utils/tinyimagenet_reformat.py:1-33): move ``val/images/<file>`` into
``val/<wnid>/<file>`` following ``val_annotations.txt``, then remove the annotation file and the empty folder, so the
validation split has the ImageFolder layout of the training split.
"""
from __future__ import annotations

import os
import shutil
import sys
from typing import List

DROPPED_COLUMNS = (
    "id member_id emp_title issue_d zip_code emp_length title earliest_cr_line last_pymnt_d hardship_start_date desc "
    "hardship_end_date payment_plan_start_date next_pymnt_d settlement_date last_credit_pull_d "
    "debt_settlement_flag_date sec_app_earliest_cr_line "
    "url mths_since_last_delinq mths_since_last_major_derog mths_since_last_record annual_inc_joint dti_joint "
    "verification_status_joint mths_since_recent_bc_dlq mths_since_recent_revol_delinq revol_bal_joint "
    "sec_app_inq_last_6mths sec_app_mort_acc sec_app_open_acc sec_app_revol_util sec_app_open_act_il "
    "sec_app_num_rev_accts sec_app_chargeoff_within_12_mths sec_app_collections_12_mths_ex_med "
    "sec_app_mths_since_last_major_derog hardship_type hardship_reason hardship_status deferral_term hardship_amount "
    "hardship_length hardship_dpd hardship_loan_status orig_projected_additional_accrued_interest "
    "hardship_payoff_balance_amount hardship_last_payment_amount settlement_status settlement_amount "
    "settlement_percentage settlement_term"
).split()


def preprocess_loan(raw_csv: str, out_dir: str = "./data/loan") -> List[str]:
    import pandas as pd
    df = pd.read_csv(raw_csv, low_memory=False)
    df = df.drop(columns=[c for c in DROPPED_COLUMNS if c in df.columns])
    # columns that are entirely empty come back as float NaN; text columns get 0 for missing like the reference
    df = df.fillna(0)
    for col in list(df.columns):
        if col == "addr_state":
            continue
        s = df[col]
        if s.dtype == object or str(s.dtype).startswith(("str", "string")):
            codes, _ = pd.factorize(s, sort=False)          # code = rank of first appearance
            df[col] = codes.astype("int64")
        elif s.dtype.kind in "fi":
            mean = float(s.mean())
            if 10.0 < mean <= 100.0:
                df[col] = s / 10
            elif 100.0 < mean <= 1000.0:
                df[col] = s / 100
            elif mean > 1000.0:
                df[col] = s / 10000
    os.makedirs(out_dir, exist_ok=True)
    paths = []
    for state, part in df.groupby("addr_state", sort=True):
        path = os.path.join(out_dir, f"loan_{state}.csv")
        part.drop(columns=["addr_state"]).to_csv(path, index=False)
        paths.append(path)
    return paths


def reformat_tiny_val(val_dir: str) -> int:
    ann = os.path.join(val_dir, "val_annotations.txt")
    img_dir = os.path.join(val_dir, "images")
    moved = 0
    with open(ann) as f:
        table = dict(line.split("\t")[:2] for line in f if line.strip())
    for name in sorted(os.listdir(img_dir)):
        wnid = table[name]
        os.makedirs(os.path.join(val_dir, wnid), exist_ok=True)
        shutil.move(os.path.join(img_dir, name), os.path.join(val_dir, wnid, name))
        moved += 1
    os.remove(ann)
    os.rmdir(img_dir)
    return moved


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] in ("-h", "--help"):
        print(__doc__)
        return 0
    cmd, rest = argv[0], argv[1:]
    if cmd == "loan":
        out = preprocess_loan(rest[0], rest[1] if len(rest) > 1 else "./data/loan")
        print(f"wrote {len(out)} per-state files")
    elif cmd == "tiny":
        print(f"moved {reformat_tiny_val(rest[0])} validation images into class folders")
    elif cmd == "synthetic-loan":
        from . import synthetic
        out = synthetic.write_loan_csvs(rest[0] if rest else "./data/loan", int(rest[1]) if len(rest) > 1 else 60000)
        print(f"wrote {len(out)} synthetic per-state files")
    else:
        print(f"unknown command {cmd!r}\n{__doc__}")
        return 2
    return 0


if __name__ == "__main__":
    sys.exit(main())
