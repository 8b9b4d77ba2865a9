"""Runs the UNMODIFIED reference (ehsan886/DBA_mod, copied verbatim to ``baseline/_ref``) through its own entry
point ``main.py --params <yaml>`` and times its rounds.  None of this framework's models / kernels / engine are on that
path.  The reference cannot run as-is on this image, so the harness supplies a compatibility shim
*around* it - nothing inside ``baseline/_ref`` is edited:

  1. ``visdom`` is not installed           -> a stub module whose ``Visdom`` object ignores every call
  2. ``matplotlib`` is not installed       -> stub ``matplotlib.pyplot``
  3. ``yaml.load(f)`` without a Loader     -> defaulted to ``FullLoader`` (PyYAML >= 6 made it mandatory)
  4. no network / datasets                 -> ``torchvision.datasets.CIFAR10 / MNIST / ImageFolder`` replaced by in-memory
                                              synthetic datasets of the same shape (the generator this framework's arm
                                              uses too)
  5. ``DataLoader(num_workers=8)``         -> ``num_workers=0`` (in-memory tensors; worker start-up would only slow the
                                              reference down)
  6. ``saved_models/`` must pre-exist      -> run inside a scratch working directory that has it
  7. FedAvg / RFA add float updates into int64 ``num_batches_tracked`` (helper.py:256,365) which modern torch rejects
     -> while those two methods run, ``Tensor.add_`` casts a floating update into an integer destination (the
     truncating behaviour of the torch <= 1.4 the code was written for)

Timing: a time stamp (after ``torch.cuda.synchronize()``) is taken every time the reference finishes a round, i.e.
each time it calls ``csv_record.save_result_csv`` (main.py:235).
"""
from __future__ import annotations

import contextlib
import io
import json
import os
import runpy
import shutil
import sys
import tempfile
import time
import types

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def reference_dir() -> str:
    """baseline/_ref (the copy ``__graft_entry__.build()`` makes; it travels to the GPU box), else the read-only
    mount the copy is made from.  '' when neither exists."""
    for d in (os.path.join(HERE, "_ref"), "/root/reference"):
        if os.path.exists(os.path.join(d, "main.py")):
            return d
    return ""


REF = reference_dir() or os.path.join(HERE, "_ref")


def _install_stubs():
    import yaml
    vis = types.ModuleType("visdom")

    class Visdom:  # noqa: D401
        def __init__(self, *a, **k):
            pass

        def __getattr__(self, name):
            def _noop(*a, **k):
                return False if name == "win_exists" else None
            return _noop

    vis.Visdom = Visdom
    sys.modules["visdom"] = vis
    if "matplotlib" not in sys.modules:
        try:
            import matplotlib.pyplot  # noqa: F401
        except Exception:  # noqa: BLE001
            mpl = types.ModuleType("matplotlib")
            plt = types.ModuleType("matplotlib.pyplot")

            def _plt_getattr(name):
                if name.startswith("__"):
                    raise AttributeError(name)
                return lambda *a, **k: None

            plt.__getattr__ = _plt_getattr  # type: ignore[attr-defined]
            mpl.pyplot = plt
            sys.modules["matplotlib"] = mpl
            sys.modules["matplotlib.pyplot"] = plt
    try:
        import cv2  # noqa: F401
    except Exception:  # noqa: BLE001
        sys.modules["cv2"] = types.ModuleType("cv2")
    _orig_load = yaml.load

    def _load(stream, Loader=None, **kw):
        return _orig_load(stream, Loader=Loader or yaml.FullLoader, **kw)

    yaml.load = _load


def _install_synthetic_datasets(train_size, test_size, data_seed=None):
    import numpy as np
    import torch
    import torch.utils.data as tud
    from torchvision import datasets

    sys.path.insert(0, ROOT)
    from dba_mod_b200.data import synthetic
    cache = {}

    def _arrays(kind):
        if kind not in cache:
            cache[kind] = synthetic.make_images(kind, train_size, test_size,
                                                seed=synthetic.DATA_SEED if data_seed is None else data_seed)
        return cache[kind]

    def _make(kind, folder=False):
        class _Synth(tud.Dataset):
            def __init__(self, root=None, transform=None, train=True, download=False, **kw):
                if folder:                       # ImageFolder(root, transform): the split is in the path
                    train = "train" in str(root)
                xtr, ytr, xte, yte = _arrays(kind)
                x, y = (xtr, ytr) if train else (xte, yte)
                self.data = x.numpy() if kind != "mnist" else x.numpy()[..., 0]
                self.targets = [int(v) for v in y.tolist()]
                self.transform = transform

            def __len__(self):
                return len(self.targets)

            def __getitem__(self, i):
                img = self.data[i]
                if img.ndim == 2:
                    img = img[:, :, None]
                if folder:                       # ImageFolder yields PIL images (RandomHorizontalFlip needs one)
                    from PIL import Image
                    img = Image.fromarray(np.ascontiguousarray(img))
                    return (self.transform(img) if self.transform is not None else img), self.targets[i]
                if self.transform is not None:
                    img = self.transform(np.ascontiguousarray(img))
                return img, self.targets[i]
        return _Synth

    datasets.CIFAR10 = _make("cifar")
    datasets.MNIST = _make("mnist")
    datasets.ImageFolder = _make("tiny-imagenet-200", folder=True)    # ./data/tiny-imagenet-200/{train,val}

    _DL = tud.DataLoader
    _orig_init = _DL.__init__

    def _init(self, *a, **k):
        k["num_workers"] = 0
        k.pop("pin_memory", None)
        _orig_init(self, *a, **k)

    _DL.__init__ = _init


def _install_int_add_guard():
    import torch
    import helper as ref_helper  # the reference's module (baseline/_ref on sys.path)
    _add_ = torch.Tensor.add_

    def _guarded_add_(self, other, *a, **k):
        if not self.dtype.is_floating_point and isinstance(other, torch.Tensor) and other.dtype.is_floating_point:
            other = other.to(self.dtype)
        return _add_(self, other, *a, **k)

    def _wrap(fn):
        def inner(*a, **k):
            torch.Tensor.add_ = _guarded_add_
            try:
                return fn(*a, **k)
            finally:
                torch.Tensor.add_ = _add_
        return inner

    ref_helper.Helper.average_shrink_models = _wrap(ref_helper.Helper.average_shrink_models)
    ref_helper.Helper.geometric_median_update = _wrap(ref_helper.Helper.geometric_median_update)


def run_reference(params: dict, warmup: int, steps: int, train_size=None, test_size=None, log_path=None,
                  data_seed=None, records: bool = False) -> dict:
    """Runs rounds 1 .. warmup+steps of the reference with ``params`` (a dict in the reference's YAML schema) and
    returns {'ms_per_round': ..., 'round_end_times': [...], 'global_acc': ..., ...}."""
    import torch
    import yaml

    if not os.path.isdir(REF) or not os.path.exists(os.path.join(REF, "main.py")):
        raise FileNotFoundError("baseline/_ref is missing (copy /root/reference there)")
    work = tempfile.mkdtemp(prefix="dba_ref_")
    os.makedirs(os.path.join(work, "saved_models"))
    os.makedirs(os.path.join(work, "utils"))
    params = dict(params)
    params["epochs"] = warmup + steps
    with open(os.path.join(work, "utils", "bench_params.yaml"), "w") as f:
        yaml.safe_dump(params, f)

    if params.get("type") == "loan":      # the reference reads ./data/loan/loan_XX.csv (loan_helper.py:117-140)
        sys.path.insert(0, ROOT)
        from dba_mod_b200.data import synthetic
        synthetic.write_loan_csvs(os.path.join(work, "data", "loan"), int(train_size or 60000),
                                  synthetic.DATA_SEED if data_seed is None else data_seed)
    _install_stubs()
    _install_synthetic_datasets(train_size, test_size, data_seed)
    sys.path.insert(0, REF)
    import utils.csv_record as csv_record  # reference's (REF first on sys.path; namespace differs from ours)
    _install_int_add_guard()

    stamps = []
    _orig_save = csv_record.save_result_csv

    def _save(*a, **k):
        r = _orig_save(*a, **k)
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        stamps.append(time.perf_counter())
        return r

    csv_record.save_result_csv = _save
    old_argv, old_cwd = sys.argv, os.getcwd()
    sys.argv = ["main.py", "--params", "utils/bench_params.yaml"]
    os.chdir(work)
    sink = open(log_path, "w") if log_path else io.StringIO()
    t_start = time.perf_counter()
    try:
        with contextlib.redirect_stdout(sink), contextlib.redirect_stderr(sink):
            import logging
            logging.getLogger("logger").handlers.clear()
            runpy.run_path(os.path.join(REF, "main.py"), run_name="__main__")
    finally:
        os.chdir(old_cwd)
        sys.argv = old_argv
        if log_path:
            sink.close()
    out = {"round_end_times": [t - t_start for t in stamps], "rounds": len(stamps)}
    if len(stamps) >= warmup + steps and steps > 0:
        t0 = stamps[warmup - 1] if warmup > 0 else t_start
        out["ms_per_round"] = (stamps[warmup + steps - 1] - t0) * 1e3 / steps
    # last global accuracy / ASR rows the reference recorded
    try:
        out["global_acc"] = [r for r in csv_record.test_result if r[0] == "global"][-1][3]
        pr = [r for r in csv_record.posiontest_result if r[0] == "global"]
        out["poison_acc"] = pr[-1][3] if pr else None
    except Exception:  # noqa: BLE001
        pass
    if records:      # everything the reference recorded, round by round (utils/csv_record.py:4-19)
        out["records"] = {"test": [list(r) for r in csv_record.test_result],
                          "poison": [list(r) for r in csv_record.posiontest_result],
                          "trigger": [list(r) for r in csv_record.poisontriggertest_result],
                          "weight": [list(r) for r in csv_record.weight_result],
                          "scale": [list(r) for r in csv_record.scale_result]}
    shutil.rmtree(work, ignore_errors=True)
    return out


if __name__ == "__main__":
    import yaml
    cfg = yaml.safe_load(open(sys.argv[1]))
    print(json.dumps(run_reference(cfg, int(sys.argv[2]), int(sys.argv[3]))))
