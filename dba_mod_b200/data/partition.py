"""Client partitions of a training set.

All three partitioners consume the SAME random streams in the SAME order as the reference, so that with equal seeds
a client holds exactly the samples it holds there (tests compare against the reference's own functions):

  * class index             -> image_helper.py:72-80   (``build_classes_dict``)
  * Dirichlet non-IID split -> image_helper.py:82-110  (python ``random.shuffle`` per class, then one numpy
                               ``dirichlet`` draw per class; a client takes ``round(p * class_size)`` samples off the
                               front of what is left of that class)
  * equal split             -> image_helper.py:231-237,265-280 (one ``random.shuffle`` of all indices, contiguous
                               slices of ``len // n_clients``; the remainder is dropped)

The functions take explicit RNG objects (``random.Random`` / ``numpy.random.RandomState``) instead of the global
modules; a ``Random(1)`` / ``RandomState(1)`` pair reproduces ``random.seed(1); np.random.seed(1)``.
"""
from __future__ import annotations

import random
from collections import defaultdict
from typing import Dict, Iterable, List

import numpy as np


def build_classes_dict(labels: Iterable[int]) -> Dict[int, List[int]]:
    """label -> sample indices in data-set order (image_helper.py:72-80, without touching the images)."""
    out: Dict[int, List[int]] = {}
    for ind, label in enumerate(labels):
        out.setdefault(int(label), []).append(ind)
    return out


def dirichlet_partition(classes: Dict[int, List[int]], n_clients: int, alpha: float, py_rng: random.Random,
                        np_rng: np.random.RandomState) -> Dict[int, List[int]]:
    """Non-IID split (image_helper.py:82-110).  ``classes`` is consumed (its lists are shuffled and emptied)."""
    class_size = len(classes[0])          # the reference sizes every class by class 0
    per_client: Dict[int, List[int]] = defaultdict(list)
    for n in range(len(classes)):
        py_rng.shuffle(classes[n])
        probs = class_size * np_rng.dirichlet(np.array(n_clients * [alpha]))
        for user in range(n_clients):
            take = min(len(classes[n]), int(round(probs[user])))
            per_client[user].extend(classes[n][:take])
            classes[n] = classes[n][take:]
    return per_client


def equal_partition(n_samples: int, n_clients: int, py_rng: random.Random) -> Dict[int, List[int]]:
    """IID split into equal contiguous slices of one shuffled index list (image_helper.py:231-237,265-280)."""
    all_range = list(range(n_samples))
    py_rng.shuffle(all_range)
    per = int(n_samples / n_clients)
    return {pos: all_range[pos * per:(pos + 1) * per] for pos in range(n_clients)}


def class_histogram(parts: Dict[int, List[int]], labels, n_classes: int) -> np.ndarray:
    """[client, class] sample counts: the numbers behind the reference's (commented-out) Dirichlet bar chart
    (image_helper.py:109,112-146) and its ``__main__`` print (image_helper.py:352-377)."""
    lab = np.asarray(labels)
    keys = sorted(parts.keys())
    hist = np.zeros((len(keys), n_classes), dtype=np.int64)
    for r, k in enumerate(keys):
        if len(parts[k]):
            hist[r] = np.bincount(lab[np.asarray(parts[k], dtype=np.int64)], minlength=n_classes)
    return hist
