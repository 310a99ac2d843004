"""Test-only stand-in for the `imageio` package (absent from this image; view_synthesis/datasets/dataset.py:6,70 needs
`imageio.imread`).  Put tests/shims on PYTHONPATH of the subprocess that runs the reference's train.py / eval.py; it is
NOT part of the product package."""
import numpy as _np
from PIL import Image as _Image


def imread(path):
    with _Image.open(path) as im:
        return _np.asarray(im)
