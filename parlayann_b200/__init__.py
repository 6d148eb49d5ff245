"""parlayann_b200 — ParlayANN's data-parallel hot path (beam search, Vamana build, exact ground truth)
on B200 (sm_100a) behind a C ABI (include/pann.h).  No CPU fallback."""
from .index import (GraphIndex, build_vamana_index, load_index, defaults, random_permutation, batch_schedule,
                    groundtruth_merge)
from ._lib import PannError, QueryParams, device_count
from . import fileio, harness, sharded

__all__ = ["GraphIndex", "build_vamana_index", "load_index", "defaults", "random_permutation", "batch_schedule",
           "groundtruth_merge", "PannError", "QueryParams", "device_count", "fileio", "harness", "sharded"]
