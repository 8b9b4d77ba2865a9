"""Per-round client selection (reference main.py:139-165).

Three modes:
  * ``is_random_namelist`` and ``is_random_adversary``: uniform ``random.sample(participants, no_models)``; whoever
    of the sample is in ``adversary_list`` counts as adversarial.
  * ``is_random_namelist`` only: adversaries whose ``{idx}_poison_epochs`` contains one of the on-going global
    epochs are forced in (in adversary_list order); the remaining seats are sampled from
    ``benign_namelist + idle adversaries`` (Q27).
  * otherwise: the fixed ``participants_namelist``; all of ``adversary_list`` counts as adversarial unless
    ``is_random_adversary``.
The python RNG object is the one that already drove the partition shuffles, which keeps the reference's stream.
"""
from __future__ import annotations

import copy
import random
from typing import List, Tuple


def select_clients(params, epoch: int, participants_list: List, benign_namelist: List,
                   py_rng: random.Random) -> Tuple[List, List]:
    agent_name_keys = list(participants_list)
    adversarial_name_keys: List = []
    if params["is_random_namelist"]:
        if params["is_random_adversary"]:
            agent_name_keys = py_rng.sample(list(participants_list), params["no_models"])
            adversarial_name_keys = [n for n in agent_name_keys if n in params["adversary_list"]]
        else:
            ongoing = range(epoch, epoch + params["aggr_epoch_interval"])
            for idx, adv in enumerate(params["adversary_list"]):
                if any(e in params.poison_epochs_of(idx) for e in ongoing) and adv not in adversarial_name_keys:
                    adversarial_name_keys.append(adv)
            idle = [copy.deepcopy(a) for a in params["adversary_list"] if a not in adversarial_name_keys]
            seats = params["no_models"] - len(adversarial_name_keys)
            agent_name_keys = adversarial_name_keys + py_rng.sample(list(benign_namelist) + idle, seats)
    elif not params["is_random_adversary"]:
        adversarial_name_keys = copy.deepcopy(list(params["adversary_list"]))
    return agent_name_keys, adversarial_name_keys
