"""CSV result recorder.

Schema parity with the reference's ``utils/csv_record.py:4-64`` (six files, same headers, same row layouts,
``weight_result.csv`` and ``scale_result.csv`` headerless).  The reference keeps module-level lists that every
module appends to and rewrites all files from scratch each round (Q26); here the state lives in a
:class:`Recorder` instance owned by the simulator (several simulations can coexist in one process, which the
tests rely on), and files are still rewritten in full so a crash never leaves a half-appended row.
"""
from __future__ import annotations

import copy
import csv
import os
from typing import List

TRAIN_HEADER = ["local_model", "round", "epoch", "internal_epoch", "average_loss", "accuracy", "correct_data",
                "total_data"]
TEST_HEADER = ["model", "epoch", "average_loss", "accuracy", "correct_data", "total_data"]
TRIGGER_HEADER = ["model", "trigger_name", "trigger_value", "epoch", "average_loss", "accuracy", "correct_data",
                  "total_data"]


class Recorder:
    def __init__(self) -> None:
        self.train_result: List[list] = []
        self.test_result: List[list] = []
        self.posiontest_result: List[list] = []          # [sic] file name kept: posiontest_result.csv
        self.poisontriggertest_result: List[list] = []
        self.weight_result: List[list] = []
        self.scale_result: List[list] = []
        self.scale_temp_one_row: List = []

    # -- appenders (mirror the call sites in main.py / image_train.py / helper.py) --------------------------
    def add_train(self, name, temp_local_epoch, epoch, internal_epoch, loss, acc, correct, total) -> None:
        self.train_result.append([name, temp_local_epoch, epoch, internal_epoch, loss, acc, correct, total])

    def add_test(self, name, epoch, loss, acc, correct, total) -> None:
        self.test_result.append([name, epoch, loss, acc, correct, total])

    def add_poison_test(self, name, epoch, loss, acc, correct, total) -> None:
        self.posiontest_result.append([name, epoch, loss, acc, correct, total])

    def add_trigger_test(self, model, trigger_name, trigger_value, epoch, loss, acc, correct, total) -> None:
        self.poisontriggertest_result.append([model, trigger_name, trigger_value, epoch, loss, acc, correct, total])

    def add_weight_result(self, names, weights, alphas) -> None:
        """Three headerless rows per round: names / weights / alphas (csv_record.py:61-64)."""
        self.weight_result.append(list(names))
        self.weight_result.append(list(weights))
        self.weight_result.append(list(alphas))

    def add_scale(self, epoch, distance) -> None:
        """image_train.py:176-177 appends [epoch, round(dist, 4)] per scaling attacker."""
        self.scale_temp_one_row.append(epoch)
        self.scale_temp_one_row.append(round(float(distance), 4))

    def close_scale_row(self, global_acc) -> None:
        """main.py:202-203: the round's global accuracy terminates the row, if any attacker scaled."""
        if len(self.scale_temp_one_row) > 0:
            self.scale_temp_one_row.append(round(float(global_acc), 4))

    def flush_scale_row(self) -> bool:
        """Move the round's pending scale row into ``scale_result`` (the reference does this inside save_result_csv,
        utils/csv_record.py:52-56); also called by runs that keep the records in memory only."""
        if len(self.scale_temp_one_row) == 0:
            return False
        self.scale_result.append(copy.deepcopy(self.scale_temp_one_row))
        self.scale_temp_one_row.clear()
        return True

    # -- writer -----------------------------------------------------------------------------------------------
    def save_result_csv(self, epoch: int, is_poison: bool, folder_path: str) -> None:
        def _write(name, header, rows):
            with open(os.path.join(folder_path, name), "w", newline="") as f:
                w = csv.writer(f)
                if header is not None:
                    w.writerow(header)
                w.writerows(rows)

        _write("train_result.csv", TRAIN_HEADER, self.train_result)
        _write("test_result.csv", TEST_HEADER, self.test_result)
        if len(self.weight_result) > 0:
            _write("weight_result.csv", None, self.weight_result)
        if self.flush_scale_row():
            _write("scale_result.csv", None, self.scale_result)
        if is_poison:
            _write("posiontest_result.csv", TEST_HEADER, self.posiontest_result)
            _write("poisontriggertest_result.csv", TRIGGER_HEADER, self.poisontriggertest_result)
