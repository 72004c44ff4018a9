"""twtml-b200: the per-micro-batch learning path of twtml-spark on NVIDIA B200.

The directory name carries a hyphen (it follows the reference's repository name), so it is not
importable by that name; `import twtml_b200` (the alias package at the repo root) loads it.
"""
from . import _lib
from ._lib import TwtmlError, build
from .api import (Batch, Context, DenseVector, DStream, LabeledPoint, LinearRegressionModel, MllibHelper,
                  SparseVector, StandardScaler, StreamingKMeans, StreamingKMeansModel, StreamingLinearRegressionWithSGD, TweetBatch,
                  Vectors, lightning_series, stats_payload)

__all__ = ["Batch", "Context", "DenseVector", "DStream", "LabeledPoint", "LinearRegressionModel", "MllibHelper",
           "SparseVector", "StandardScaler", "StreamingKMeans", "StreamingKMeansModel", "StreamingLinearRegressionWithSGD", "TweetBatch", "TwtmlError", "Vectors", "build", "lightning_series", "stats_payload", "_lib"]
