"""Reference-compatible entry point: ``from model import Model`` (reference: model.py)."""
from tensorflowdistributedlearning_b200.api import (BASE_DEPTH, BATCH_NORM_DECAY, BATCH_NORM_EPSILON,  # noqa: F401
                                                    BATCH_NORM_SCALE, INPUT_SHAPE, OUTPUT_STRIDE, WEIGHT_DECAY, Model)
