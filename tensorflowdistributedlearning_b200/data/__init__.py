"""Input pipeline: decode/augment (CPU oracle + CUDA warp), K-fold directories, datasets, batch loader."""
from . import folds, preprocessing  # noqa: F401
from .dataset import BatchLoader, SegmentationFolder, SyntheticClassification, SyntheticSegmentation  # noqa: F401
