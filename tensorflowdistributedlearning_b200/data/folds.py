"""K-fold directory management (reference: preprocessing/preprocessing.py:33-88, model.py:134-186).

Layout kept from the reference so existing model directories stay usable:
``{model_dir}/{train,eval}/{images,masks}/fold{k}/{id}.png`` symlinks into ``{data_dir}/{images,masks}/{id}.png``.
Differences (SURVEY defects 7, 10): fold directories are only wiped when asked (``clean=True``), the
"already linked" guard actually looks inside the fold directory, and image/mask pairing is by id, not by
the order of two independent globs.
"""
from __future__ import annotations

import glob
import os
from typing import Iterable, List, Sequence, Tuple

import numpy as np

MODES = ("train", "eval")
KINDS = ("images", "masks")


def prepare_directory(model_directory: str, n_folds: int = 5, clean: bool = False) -> None:
    """Create the fold directory tree; with ``clean`` delete the files (links) already inside, like the reference."""
    for mode in MODES:
        for kind in KINDS:
            for k in range(n_folds):
                d = os.path.join(model_directory, mode, kind, f"fold{k}")
                os.makedirs(d, exist_ok=True)
                if clean:
                    for f in os.listdir(d):
                        p = os.path.join(d, f)
                        if os.path.islink(p) or os.path.isfile(p):
                            os.unlink(p)


_prepare_directory = prepare_directory  # reference name


def create_symlinks(data_dir: str, model_dir: str, mode: str, idx: Iterable[str], fold: int) -> int:
    """Link ``{data_dir}/{images,masks}/{id}.png`` into the fold directories.  Returns the number of new links."""
    if mode not in MODES:
        raise ValueError(f"mode must be one of {MODES}")
    made = 0
    for kind in KINDS:
        d = os.path.join(model_dir, mode, kind, f"fold{fold}")
        os.makedirs(d, exist_ok=True)
        for i in idx:
            src = os.path.abspath(os.path.join(data_dir, kind, f"{i}.png"))
            dst = os.path.join(d, f"{i}.png")
            if os.path.lexists(dst):
                continue
            os.symlink(src, dst)
            made += 1
    return made


def list_fold(model_dir: str, mode: str, fold: int) -> List[Tuple[str, str]]:
    """(image, mask) path pairs of a fold, paired by id and sorted."""
    img_dir = os.path.join(model_dir, mode, "images", f"fold{fold}")
    msk_dir = os.path.join(model_dir, mode, "masks", f"fold{fold}")
    pairs = []
    for p in sorted(glob.glob(os.path.join(img_dir, "*.png"))):
        m = os.path.join(msk_dir, os.path.basename(p))
        if not os.path.lexists(m):
            raise FileNotFoundError(f"mask missing for {p}")
        pairs.append((p, m))
    return pairs


def coverage_class(mask: np.ndarray) -> int:
    """Driver recipe (Test.ipynb:111-118): smallest i in [0,10] with coverage*10 <= i."""
    cov = float(mask.sum()) / float(mask.size)
    for i in range(11):
        if cov * 10 <= i:
            return i
    return 10


def stratified_folds(X: Sequence, y: Sequence, n_splits: int = 5, seed: int = 42):
    """sklearn StratifiedKFold(n_splits, shuffle=True, random_state=seed) splits (reference: model.py:134-154)."""
    from sklearn.model_selection import StratifiedKFold

    skf = StratifiedKFold(n_splits=n_splits, shuffle=True, random_state=seed)
    return list(skf.split(np.asarray(X), np.asarray(y)))
