"""Row N1 of the coverage table: the reference's own train.py and eval.py, UNCHANGED, against the drop-in package.

The reference tree is the byte-for-byte copy under oracle/_ref/ (oracle/make_ref.py; verified against its manifest).  Each
driver runs in a subprocess with

    PYTHONSAFEPATH = 1    PYTHONPATH = code-nerf_b200 : tests/shims (imageio stand-in) : oracle/_ref
    CODENERF_REFERENCE = oracle/_ref

(PYTHONSAFEPATH keeps the interpreter from putting the script's own directory -- the reference checkout, which holds a
`view_synthesis` package of its own -- in front of PYTHONPATH; without it `python <ref>/train.py` silently imports the
reference's kernels-free package and the drop-in is never touched.  The launch-count check below is what catches that.)

so `view_synthesis.nerf` / `view_synthesis.models` resolve to the drop-in and `view_synthesis.utils / datasets / cfgnode`,
train.py and eval.py to the reference (train.py:12-15, eval.py:15-18).  Data: a tiny SRN-layout tree written by
scripts/make_synthetic_srn.py; configs: the reference's srn-cars-code.yml / srn-cars-code-3080-val.yml with only paths and
iteration counts edited.  Checked: exit code 0, a checkpoint file, finite and decreasing training loss, the validation
render inside train.py (eval.validate: test-time optimisation of codes + camera pose, then parallel_image_render), and
that libcnrf kernels were launched by the driver process (CODENERF_REPORT_LAUNCHES).
"""
import importlib.util
import os
import re
import subprocess
import sys

import pytest
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
pytestmark = pytest.mark.gpu


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _env():
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([os.path.join(ROOT, "code-nerf_b200"), os.path.join(ROOT, "tests", "shims"), REF])
    env["CODENERF_REFERENCE"] = REF
    env["PYTHONSAFEPATH"] = "1"
    env["CODENERF_REPORT_LAUNCHES"] = "1"
    env["PYTHONDONTWRITEBYTECODE"] = "1"
    return env


def _run(args, cwd, timeout=900):
    p = subprocess.run([sys.executable] + args, cwd=cwd, env=_env(), stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                       text=True, timeout=timeout)
    print(p.stdout[-6000:])
    return p


def _launches(out):
    return [int(m) for m in re.findall(r"\[libcnrf\] pid \d+: (\d+) kernel launches", out)]


def _field(out, tag, key):
    return [float(m) for m in re.findall(rf"\[{tag}\s*\].*?{key}: ([-0-9.einfa]+)", out)]


@pytest.fixture(scope="module")
def workdir(tmp_path_factory):
    assert _load(os.path.join(ROOT, "oracle", "make_ref.py"), "make_ref").verify(), \
        "oracle/_ref is missing or edited: run `python oracle/make_ref.py` in the build container"
    wd = tmp_path_factory.mktemp("n1")
    data = _load(os.path.join(ROOT, "scripts", "make_synthetic_srn.py"), "make_srn").main(str(wd / "data"), 6, 2, 10, 128)
    return wd, data


def _train_cfg(wd, data, name, iterations=96):
    cfg = yaml.safe_load(open(os.path.join(REF, "config", "srn-cars-code.yml")))
    cfg["experiment"].update(id=name, logdir=str(wd / "logs"), iterations=iterations, print_every=8, save_every=40,
                             validate_every=64, val_iterations=6, val_print_every=2)
    cfg["dataset"]["basedir"] = data
    # eval.validate (called from train.py:142) reads these two; srn-cars-code.yml itself does not carry them
    cfg["optimizer"].update(angle_lr=5.0e-2, radius_lr=1.0e-1)
    path = wd / f"{name}.yml"
    path.write_text(yaml.safe_dump(cfg))
    return str(path)


@pytest.fixture(scope="module")
def trained(workdir):
    wd, data = workdir
    p = _run([os.path.join(REF, "train.py"), "-c", _train_cfg(wd, data, "n1-train")], cwd=str(wd))
    return wd, data, p


def test_reference_train_py_runs_unchanged_on_the_drop_in(trained):
    wd, _, p = trained
    assert p.returncode == 0, p.stdout[-3000:]
    ckpts = sorted((wd / "logs" / "n1-train").glob("*.ckpt"))
    assert ckpts, "train.py wrote no checkpoint"
    fine = _field(p.stdout, "TRAIN", "nerf_loss_fine")
    assert len(fine) >= 8 and all(f == f and f < 1e3 for f in fine), fine
    assert min(fine[-3:]) < fine[0], f"fine loss did not decrease: {fine}"
    # the validation pass inside train.py: test-time optimisation lines and the rendered view
    assert _field(p.stdout, "VALOPT", "psnr"), "eval.validate did not run its optimisation loop"
    assert _field(p.stdout, "VAL", "psnr"), "parallel_image_render did not produce a view"
    n = _launches(p.stdout)
    assert n and max(n) > 1000, f"the driver process launched no libcnrf kernels: {n}"
    ck = torch.load(str(ckpts[-1]), map_location="cpu")
    assert sorted(ck) == ["iter", "model_embedding_state_dict", "model_nerf_coarse_state_dict",
                          "model_nerf_fine_state_dict", "optimizer_state_dict"]
    assert ck["model_nerf_fine_state_dict"]["layer_xyz1.weight"].shape == (256, 63)


def test_reference_eval_py_runs_unchanged_on_the_drop_in(trained):
    wd, data, ptrain = trained
    assert ptrain.returncode == 0
    ckpt = sorted((wd / "logs" / "n1-train").glob("*.ckpt"))[-1]
    cfg = yaml.safe_load(open(os.path.join(REF, "config", "srn-cars-code-3080-val.yml")))
    # eval.validate skips five loader batches (eval.py:109) and the "val" sampler draws `iterations` samples
    # (utils/util.py:72-76): at least six; every iteration is one validate() call = one optimisation + one render
    cfg["experiment"].update(id="n1-eval", logdir=str(wd / "logs"), iterations=6, val_iterations=12, val_print_every=4)
    cfg["dataset"]["basedir"] = data
    path = wd / "n1-eval.yml"
    path.write_text(yaml.safe_dump(cfg))
    p = _run([os.path.join(REF, "eval.py"), "-c", str(path), "--load-checkpoint", str(ckpt)], cwd=str(wd))
    assert p.returncode == 0, p.stdout[-3000:]
    total = _field(p.stdout, "VALOPT", "total_loss")
    per = len(total) // 6
    assert per >= 2 and all(t == t for t in total), total
    # six validate() calls: within each the optimised loss must not end above where it started
    assert sum(total[k * per + per - 1] < total[k * per] for k in range(6)) >= 5, total
    psnr = _field(p.stdout, "VAL", "psnr")
    assert len(psnr) == 6 and all(x == x and x > 3.0 for x in psnr), psnr
    n = _launches(p.stdout)
    assert n and max(n) > 1000, n


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_reference_train_py_distributed_two_gpus(workdir):
    """train.py --distributed -g 2: mp.spawn + NCCL + the reference's three DistributedDataParallel wrappers
    (utils/util.py:139-142) around the drop-in modules, DistributedSampler, broadcast_object_list in validate."""
    wd, data = workdir
    p = _run([os.path.join(REF, "train.py"), "-c", _train_cfg(wd, data, "n1-train-ddp", iterations=96), "--distributed", "-g", "2"],
             cwd=str(wd))
    assert p.returncode == 0, p.stdout[-3000:]
    fine = _field(p.stdout, "TRAIN", "nerf_loss_fine")
    assert len(fine) >= 8 and min(fine[-3:]) < fine[0], fine
    assert _field(p.stdout, "VAL", "psnr"), "distributed parallel_image_render did not produce a view"
    n = _launches(p.stdout)
    assert len(n) >= 2 and min(n) > 1000, f"both ranks must launch libcnrf kernels: {n}"
    assert sorted((wd / "logs" / "n1-train-ddp").glob("*.ckpt"))
