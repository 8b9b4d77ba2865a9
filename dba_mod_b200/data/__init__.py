"""Data layer: device-resident data sets, client partitions, trigger tables, synthetic generators, offline prep.

Reference counterparts: image_helper.py:72-350 (ImageHelper), loan_helper.py:29-210 (LoanHelper / StateHelper /
LoanDataset), utils/loan_preprocess.py, utils/tinyimagenet_reformat.py.
"""
from . import partition, prep, synthetic, triggers
from .image import ImageData
from .loan import LoanData

__all__ = ["ImageData", "LoanData", "partition", "prep", "synthetic", "triggers"]
