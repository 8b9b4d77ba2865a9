"""Data layer: device-resident data sets, client partitions, trigger tables, synthetic generators, offline prep.

Reference counterparts: image_helper.py:72-350 (ImageHelper), loan_helper.py:29-210 (LoanHelper / StateHelper /
LoanDataset), utils/loan_preprocess.py, utils/tinyimagenet_reformat.py.
"""
from . import partition, synthetic, triggers
from .image import ImageData
from .loan import LoanData

__all__ = ["ImageData", "LoanData", "partition", "prep", "synthetic", "triggers"]


def __getattr__(name):
    # `prep` is also an entry point (python -m dba_mod_b200.data.prep): importing it here eagerly makes runpy warn
    if name == "prep":
        import importlib
        return importlib.import_module(".prep", __name__)
    raise AttributeError(name)
