"""Row f4 (data path), host side: the drop-in `view_synthesis.datasets.SRNDataset` against the reference's own class
(oracle/_ref copy, imageio stand-in from tests/shims) on a synthetic SRN tree -- every sample bit-identical, same
attributes and printed lines, same batches through the reference's DataLoader recipe (utils/util.py:72-88)."""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="module")
def trees(tmp_path_factory):
    assert os.path.isfile(os.path.join(REF, "view_synthesis", "datasets", "dataset.py")), \
        "oracle/_ref missing: run `python oracle/make_ref.py` in the build container"
    root = tmp_path_factory.mktemp("srn")
    data = _load(os.path.join(ROOT, "scripts", "make_synthetic_srn.py"), "make_srn").main(str(root), 3, 2, 4, 128)
    sys.path.insert(0, os.path.join(ROOT, "tests", "shims"))          # imageio stand-in for the reference module
    try:
        ref_mod = _load(os.path.join(REF, "view_synthesis", "datasets", "dataset.py"), "ref_dataset")
    finally:
        sys.path.pop(0)
    ours_mod = _load(os.path.join(ROOT, "code-nerf_b200", "view_synthesis", "datasets", "dataset.py"), "dropin_dataset")
    return data, ref_mod, ours_mod


@pytest.mark.parametrize("stage", ["train", "val"])
def test_drop_in_dataset_samples_are_bit_identical(trees, stage, capsys):
    data, ref_mod, ours_mod = trees
    ref = ref_mod.SRNDataset(data, stage)
    ref_out = capsys.readouterr().out
    ours = ours_mod.SRNDataset(data, stage)
    assert capsys.readouterr().out == ref_out                        # the three printed lines
    assert len(ours) == len(ref) and ours.num_objects == ref.num_objects and ours.num_views == ref.num_views
    assert ours.dataset_name == ref.dataset_name and ours.base_path == ref.base_path and ours.stage == ref.stage
    assert ours.rgb_all_filenames == ref.rgb_all_filenames and ours.pose_all_filenames == ref.pose_all_filenames
    for rep in range(2):                                             # second pass: served from the cache
        for i in range(len(ref)):
            a, b = ref[i], ours[i]
            assert sorted(a) == sorted(b)
            assert a["object_id"] == b["object_id"]
            for k in ("intrinsic", "color", "mask", "pose"):
                assert a[k].dtype == b[k].dtype == np.float32 and a[k].shape == b[k].shape, k
                assert np.array_equal(a[k], b[k]), (k, i, rep)
    s = ours[0]
    s["color"][:] = -1.0; s["pose"][:] = 7.0                         # a caller scribbling on a sample must not reach the cache
    assert np.array_equal(ours[0]["color"], ref[0]["color"]) and np.array_equal(ours[0]["pose"], ref[0]["pose"])


def test_drop_in_dataset_without_cache_and_through_the_reference_loader_recipe(trees, monkeypatch):
    data, ref_mod, ours_mod = trees
    monkeypatch.setenv("CODENERF_DATASET_CACHE", "0")
    nocache = ours_mod.SRNDataset(data, "train")
    monkeypatch.delenv("CODENERF_DATASET_CACHE")
    cached, ref = ours_mod.SRNDataset(data, "train"), ref_mod.SRNDataset(data, "train")
    assert nocache[3]["color"].tobytes() == ref[3]["color"].tobytes() and not nocache._views

    def batches(ds):                                                  # utils/util.py:72-88 with a fixed generator
        g = torch.Generator().manual_seed(5)
        sampler = torch.utils.data.RandomSampler(ds, replacement=True, num_samples=6, generator=g)
        dl = torch.utils.data.DataLoader(ds, batch_size=2, shuffle=False, num_workers=0, sampler=sampler, pin_memory=False)
        return list(dl)

    for a, b in zip(batches(ref), batches(cached)):
        for k in a:
            assert torch.equal(a[k], b[k]), k


def test_package_resolution_prefers_the_drop_in_dataset():
    """With the drop-in ahead of the reference on sys.path, `view_synthesis.datasets` is the cached class while
    `view_synthesis.utils` / `cfgnode` still come from the reference (view_synthesis/__init__.py extends __path__)."""
    import subprocess
    code = ("import view_synthesis.datasets as d, view_synthesis.utils as u\n"
            "print(d.SRNDataset.__module__, '|', d.__file__)\nprint(u.__file__)\n")
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([os.path.join(ROOT, "code-nerf_b200"), os.path.join(ROOT, "tests", "shims"), REF]),
               CODENERF_REFERENCE=REF, PYTHONDONTWRITEBYTECODE="1")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True).stdout.splitlines()
    assert "code-nerf_b200" in out[0] and hasattr(__import__("numpy"), "array")
    assert out[1].startswith(REF)
