"""Constants and the experiment-parameter schema.

Parity targets in the reference (read-only at /root/reference):
  * constants           -> config.py:1-13
  * YAML schema         -> utils/{mnist,cifar,tiny,loan}_params.yaml
  * dynamic numbered keys ``{i}_poison_pattern`` / ``{i}_poison_epochs`` /
    ``{i}_poison_trigger_names|values`` -> image_helper.py:328-350, loan_train.py:47-57

The reference passes a raw mutable dict around.  Here the dict is kept (the
YAML files and ``params.yaml`` dump stay byte-compatible in schema) but it is
wrapped in :class:`Params`, which validates types, fills defaults for keys the
reference ``.get()``s, and gives typed accessors for the numbered keys.
"""
from __future__ import annotations

import copy
from typing import Any, Dict, Iterable, List, Sequence

import yaml

AGGR_MEAN = "mean"
AGGR_GEO_MED = "geom_median"
AGGR_FOOLSGOLD = "foolsgold"
AGGREGATORS = (AGGR_MEAN, AGGR_GEO_MED, AGGR_FOOLSGOLD)

TYPE_LOAN = "loan"
TYPE_CIFAR = "cifar"
TYPE_MNIST = "mnist"
TYPE_TINYIMAGENET = "tiny-imagenet-200"
IMAGE_TYPES = (TYPE_CIFAR, TYPE_MNIST, TYPE_TINYIMAGENET)
ALL_TYPES = IMAGE_TYPES + (TYPE_LOAN,)

# Present in the reference's config.py but never read by it (config.py:7-8).
MAX_UPDATE_NORM = 1000
patience_iter = 20

DEFAULT_NAME = {TYPE_LOAN: "loan", TYPE_CIFAR: "cifar", TYPE_MNIST: "mnist", TYPE_TINYIMAGENET: "tiny"}

# key -> (python type(s), required)
_REQUIRED: Dict[str, tuple] = {
    "type": (str,),
    "lr": (int, float),
    "momentum": (int, float),
    "decay": (int, float),
    "batch_size": (int,),
    "test_batch_size": (int,),
    "epochs": (int,),
    "internal_epochs": (int,),
    "aggr_epoch_interval": (int,),
    "aggregation_methods": (str,),
    "no_models": (int,),
    "number_of_total_participants": (int,),
    "is_random_namelist": (bool,),
    "is_random_adversary": (bool,),
    "is_poison": (bool,),
    "eta": (int, float),
}

_DEFAULTS: Dict[str, Any] = {
    "participants_namelist": [],
    "adversary_list": [],
    "poison_epochs": [],
    "trigger_num": 0,
    "poison_label_swap": 0,
    "poisoning_per_batch": 0,
    "poison_lr": None,  # falls back to lr
    "internal_poison_epochs": 1,
    "poison_step_lr": False,
    "scale_weights_poison": 1.0,
    "baseline": False,
    "alpha_loss": 1.0,
    "centralized_test_trigger": False,
    "geom_median_maxiter": 10,
    "fg_use_memory": False,
    "sampling_dirichlet": False,
    "dirichlet_alpha": 0.9,
    "diff_privacy": False,
    "sigma": 0.01,
    "save_model": False,
    "save_on_epochs": [],
    "resumed_model": False,
    "resumed_model_name": "",
    "vis_train": False,
    "vis_train_batch_loss": False,
    "vis_trigger_split_test": False,
    "batch_track_distance": False,
    # ---- extensions of this framework (absent from the reference schema) ----
    "resume_required": False,     # resumed_model: true with a missing checkpoint file is an error (else: warn, start fresh)
    "synthetic_data": None,       # None = auto (use real data when found under data_dir, else synthetic)
    "synthetic_seed": None,       # seed of the synthetic data generator (None = data.synthetic.DATA_SEED)
    "synthetic_train_size": None,
    "synthetic_test_size": None,
    "data_dir": "./data",
    "seed": 1,                    # main.py:36-38,86 hard-code 1
    "client_eval": True,          # Q16: per-client full-test-set evaluation each round
    "overlap_client_eval": False,  # with overlap_global_eval: defer the per-client tests too (measured: no net gain)
    "overlap_global_eval": False,  # enqueue a round's global evaluations on a side stream, read them after the next
                                  # round's local training (records keep their order; log lines come one phase late)
    "backend": "auto",            # auto | torch | cuda
    # cuda backend, ResNet executors: 'tf32' = fp32 activations + tcgen05 kind::tf32 (the reference's precision class:
    # fp32 tensors, TF32 convolutions), 'bf16' = bf16 activations + kind::f16 (fast mode, half the bytes)
    "compute_dtype": "tf32",
    "replicas_per_gpu": 0,        # 0 = all clients of a rank in one batched group
    "n_gpus": 0,                  # 0 = all visible
    "fused_collectives": True,    # False = NCCL baseline path for the aggregation collectives
    "multistep_lr_compat": "torch2",  # Q4: 'torch2' (decay only at integral milestones) | 'bisect'
    "native_net": True,           # cuda backend: use the native tensor-core executor where it covers the model
    "native_train": True,
    "native_eval_chunk": 1000,    # upper bound on test images per evaluation pass (per model / input set)
    "cuda_graphs": True,          # replay the captured training step instead of ~125 launches per step
    "native_s2d": True,           # stride-2 convolutions from space-to-depth operands (output pixels only)
    "nvls": "auto",               # fused collectives through NVSwitch multicast (multimem): auto = from 4 ranks on
    "overlap_wgrad": True,        # weight-gradient kernels on a side stream / parallel graph branch
    "fuse_bn_sums": False,        # BatchNorm-backward sums in the epilogue of the producing input-gradient convolution
                                  # (measured SLOWER: the epilogue's extra loads cost more than the reduce kernel)
    "prewarm": True,              # allocate slabs / capture step graphs for every batch width at start-up
    "compact_tail": True,         # clients that finished their local steps stop being computed (native trainer)
}

# Keys that exist in the shipped YAML files but are never read by the reference (App. A).
IGNORED_KEYS = frozenset(
    """internal_poison_clean_epochs track_distance log_interval poison_momentum poison_decay results_json
    resumed_model_path s_norm report_train_loss report_test_loss report_poison_loss track_clusters modify_poison
    poison_type poison_image_id poison_image_id_2 size_of_secret_dataset poison_test_repeat random_compromise
    noise_level tied scale_weights bptt environment_name name current_time folder_path""".split()
)


class ParamsError(ValueError):
    pass


class Params(dict):
    """Validated flat parameter dict (still a ``dict`` so ``params['lr']`` etc. work)."""

    def __init__(self, raw: Dict[str, Any]):
        super().__init__(copy.deepcopy(raw))
        for k, v in _DEFAULTS.items():
            self.setdefault(k, copy.deepcopy(v))
        self._validate()

    # ------------------------------------------------------------------ io
    @classmethod
    def from_yaml(cls, path: str) -> "Params":
        with open(path, "r") as f:
            raw = yaml.safe_load(f)  # Q2: the reference's bare yaml.load(f) fails on PyYAML>=6
        if not isinstance(raw, dict):
            raise ParamsError(f"{path}: top level of the YAML must be a mapping")
        return cls(raw)

    def dump_yaml(self, path: str) -> None:
        """``params.yaml`` written at start-up (main.py:129-130)."""
        with open(path, "w") as f:
            yaml.safe_dump(_plain(dict(self)), f)

    # ---------------------------------------------------------- validation
    def _validate(self) -> None:
        for key, types in _REQUIRED.items():
            if key not in self:
                raise ParamsError(f"missing required parameter '{key}'")
            v = self[key]
            if bool in types:
                if not isinstance(v, bool):
                    raise ParamsError(f"'{key}' must be a bool, got {v!r}")
            elif isinstance(v, bool) or not isinstance(v, types):
                raise ParamsError(f"'{key}' must be {types}, got {v!r}")
        if self["type"] not in ALL_TYPES:
            raise ParamsError(f"unknown type {self['type']!r}; expected one of {ALL_TYPES}")
        if self["compute_dtype"] not in ("tf32", "bf16"):
            raise ParamsError(f"compute_dtype must be 'tf32' or 'bf16', got {self['compute_dtype']!r}")
        if self["aggregation_methods"] not in AGGREGATORS:
            raise ParamsError(f"unknown aggregation_methods {self['aggregation_methods']!r}")
        if self["poison_lr"] is None:
            self["poison_lr"] = self["lr"]
        if self["no_models"] <= 0 or self["batch_size"] <= 0:
            raise ParamsError("no_models and batch_size must be positive")
        if self["aggr_epoch_interval"] < 1:
            raise ParamsError("aggr_epoch_interval must be >= 1")
        if self["is_poison"]:
            adv = self["adversary_list"]
            if not isinstance(adv, list):
                raise ParamsError("adversary_list must be a list")
            if self.is_image:
                for j in range(int(self["trigger_num"])):
                    pat = self.get(f"{j}_poison_pattern")
                    if pat is None:
                        raise ParamsError(f"missing '{j}_poison_pattern' (trigger_num={self['trigger_num']})")
                    for pos in pat:
                        if len(pos) != 2:
                            raise ParamsError(f"'{j}_poison_pattern' entries must be [row, col]")
            else:
                for j in range(int(self["trigger_num"])):
                    n = self.get(f"{j}_poison_trigger_names")
                    v = self.get(f"{j}_poison_trigger_values")
                    if n is None or v is None or len(n) != len(v):
                        raise ParamsError(f"'{j}_poison_trigger_names/values' missing or of unequal length")
        if not self["is_random_namelist"] and len(self["participants_namelist"]) < self["no_models"]:
            raise ParamsError("participants_namelist shorter than no_models with is_random_namelist=false")

    # ------------------------------------------------------------ accessors
    @property
    def is_image(self) -> bool:
        return self["type"] in IMAGE_TYPES

    @property
    def is_centralized_attack(self) -> bool:
        """adversary_list of length 1 => centralized attack with the global trigger (image_train.py:47-48)."""
        return len(self["adversary_list"]) == 1

    def adversary_index(self, name) -> int:
        """Index into adversary_list; -1 if not an adversary.

        Q17: the image path compares ``int(agent) == adversary_list[i]``, LOAN compares strings.
        """
        for i, a in enumerate(self["adversary_list"]):
            if self.is_image:
                try:
                    if int(name) == a:
                        return i
                except (TypeError, ValueError):
                    continue
            elif name == a:
                return i
        return -1

    def poison_epochs_of(self, adv_index: int) -> List[int]:
        """``{i}_poison_epochs`` with the ``poison_epochs`` fallback list (image_train.py:38-44)."""
        if adv_index < 0:
            return list(self["poison_epochs"])
        return list(self.get(f"{adv_index}_poison_epochs", self["poison_epochs"]))

    def pixel_pattern(self, adv_index: int) -> List[Sequence[int]]:
        """Pixel list of one local trigger; ``-1`` = union of all ``trigger_num`` patterns."""
        if adv_index == -1:
            out: List[Sequence[int]] = []
            for j in range(int(self["trigger_num"])):
                out += list(self[f"{j}_poison_pattern"])
            return out
        return list(self[f"{adv_index}_poison_pattern"])

    def feature_trigger(self, adv_index: int):
        """(names, values) of one LOAN trigger; ``-1`` = concatenation of all (loan_train.py:47-57)."""
        if adv_index == -1:
            names, values = [], []
            for j in range(int(self["trigger_num"])):
                names += list(self[f"{j}_poison_trigger_names"])
                values += list(self[f"{j}_poison_trigger_values"])
            return names, values
        return list(self[f"{adv_index}_poison_trigger_names"]), list(self[f"{adv_index}_poison_trigger_values"])

    def unknown_keys(self) -> List[str]:
        known = set(_REQUIRED) | set(_DEFAULTS) | IGNORED_KEYS
        out = []
        for k in self:
            if k in known:
                continue
            head, _, tail = str(k).partition("_")
            if head.isdigit() and tail in (
                "poison_pattern", "poison_epochs", "poison_trigger_names", "poison_trigger_values"):
                continue
            out.append(k)
        return out


def _plain(o):
    if isinstance(o, dict):
        return {str(k): _plain(v) for k, v in o.items()}
    if isinstance(o, (list, tuple)):
        return [_plain(v) for v in o]
    if hasattr(o, "item") and not isinstance(o, (str, bytes)):
        try:
            return o.item()
        except Exception:  # noqa: BLE001
            return str(o)
    return o


def multistep_lr_factors(n_epochs: int, gamma: float = 0.1, compat: str = "torch2",
                         step_first: bool = False) -> List[float]:
    """LR multiplier in effect during internal epoch e (0-based) under the reference's
    ``MultiStepLR(milestones=[0.2*N, 0.8*N], gamma)`` stepped once per internal epoch.

    Images step the scheduler at the END of each internal epoch (image_train.py:118-120); LOAN steps it at the
    START (loan_train.py:90-92, Q18) -> ``step_first=True`` shifts the schedule by one.

    Q4: with float milestones torch>=2 keeps them in a Counter keyed by float, so the decay fires only when
    ``last_epoch == milestone`` holds exactly (N=10 -> epochs 2 and 8; N=6 or 3 -> never).  ``compat='bisect'``
    gives the torch<=1.x behaviour (``gamma ** bisect_right(milestones, epoch)``).
    """
    import bisect

    ms = sorted([0.2 * n_epochs, 0.8 * n_epochs])
    total = n_epochs + (1 if step_first else 0)
    seq = []
    f = 1.0
    for e in range(total):
        if compat == "bisect":
            f = gamma ** bisect.bisect_right(ms, e)
        elif e > 0:
            f *= gamma ** sum(1 for m in ms if m == e)
        seq.append(f)
    return seq[1:] if step_first else seq
