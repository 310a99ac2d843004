"""Row f4 (data path), device side: `codenerf_b200.data.DeviceSRN` (views resident in HBM as uint8, csrc/data.cu) against
the reference's loader arithmetic -- the collated DataLoader batch of the reference's own SRNDataset and the target gather
of train.py:77-80 -- bit for bit."""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
pytestmark = pytest.mark.gpu


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="module")
def data(tmp_path_factory):
    root = tmp_path_factory.mktemp("srn")
    path = _load(os.path.join(ROOT, "scripts", "make_synthetic_srn.py"), "make_srn").main(str(root), 4, 1, 5, 128)
    sys.path.insert(0, os.path.join(ROOT, "tests", "shims"))
    try:
        ref_mod = _load(os.path.join(REF, "view_synthesis", "datasets", "dataset.py"), "ref_dataset")
    finally:
        sys.path.pop(0)
    ours_mod = _load(os.path.join(ROOT, "code-nerf_b200", "view_synthesis", "datasets", "dataset.py"), "dropin_dataset")
    return ref_mod.SRNDataset(path, "train"), ours_mod.SRNDataset(path, "train")


@pytest.mark.parametrize("source", ["drop_in", "reference_class"])
def test_device_batch_equals_the_reference_loader_batch(data, source):
    from codenerf_b200.data import DeviceSRN
    ref, ours = data
    dev = torch.device("cuda", 0)
    dd = DeviceSRN(ours if source == "drop_in" else ref, dev)
    assert len(dd) == len(ref) == 20 and dd.colors_u8.dtype == torch.uint8
    idx = [7, 0, 19, 7]
    want = torch.utils.data.default_collate([ref[i] for i in idx])          # what DataLoader(batch_size=4) yields
    got = dd.batch(idx)
    assert sorted(got) == sorted(want)
    for k in want:
        assert got[k].shape == want[k].shape and got[k].dtype == want[k].dtype, k
        assert torch.equal(got[k].cpu(), want[k]), k


def test_device_targets_equal_train_py_gather(data):
    from codenerf_b200.data import DeviceSRN
    ref, ours = data
    dev = torch.device("cuda", 0)
    dd = DeviceSRN(ours, dev)
    rng = np.random.default_rng(3)
    idx = [3, 11, 3, 18]
    H, W = dd.height, dd.width
    for R in (1, 333, 4096):
        sel = np.stack([rng.permutation(H * W)[:R] for _ in idx])         # ray_sampler.py:76-78
        color = torch.stack([torch.from_numpy(ref[i]["color"]) for i in idx]).flatten(1, 2)          # train.py:77
        want = torch.cat([color[k, torch.from_numpy(sel[k]), :] for k in range(len(idx))], dim=0)       # :79-80
        got = dd.targets(idx, sel)
        assert got.shape == want.shape and torch.equal(got.cpu(), want)
        got3 = dd.targets(torch.tensor(idx), torch.from_numpy(sel).to(dev), channels=3)
        assert torch.equal(got3.cpu(), want[:, :3])
    dd.check()
    bad = sel.copy(); bad[1, 0] = H * W
    dd.targets(idx, bad)
    with pytest.raises(IndexError):
        dd.check()
    with pytest.raises(ValueError):
        dd.targets(idx[:2], sel)


def test_u8_to_unit_is_the_loaders_division_for_every_byte_value_and_ragged_lengths():
    import ctypes as C
    from codenerf_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for n in (0, 1, 3, 4, 255, 256, 1021):
        u8 = torch.arange(n, dtype=torch.int64).remainder(256).to(torch.uint8)
        want = torch.from_numpy((u8.numpy() / 255.0).astype(np.float32))
        src = u8.to(dev)
        out = torch.full((max(n, 1),), -1.0, device=dev)
        if n:
            _lib.check(lib.cnrf_u8_to_unit(C.c_void_p(src.data_ptr()), n, C.c_void_p(out.data_ptr()), st), "u8")
            assert torch.equal(out[:n].cpu(), want)
        # misaligned views take the scalar path
        if n > 5:
            _lib.check(lib.cnrf_u8_to_unit(C.c_void_p(src.data_ptr() + 1), n - 1, C.c_void_p(out.data_ptr()), st), "u8")
            assert torch.equal(out[:n - 1].cpu(), want[1:])
