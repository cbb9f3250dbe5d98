"""Reference-compatible ``preprocessing.preprocessing`` (reference: preprocessing/preprocessing.py)."""
from tensorflowdistributedlearning_b200.data.folds import _prepare_directory, create_symlinks  # noqa: F401
from tensorflowdistributedlearning_b200.data.preprocessing import (MEAN, STD, laplace, parse_image as _parse_image,  # noqa: F401
                                                                   read_and_preprocess, read_image,
                                                                   single_transformation,
                                                                   single_transformation_from_jpeg,
                                                                   single_transformation_from_matrix)
