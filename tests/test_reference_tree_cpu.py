"""The reference copy under oracle/_ref (oracle/make_ref.py) and the synthetic SRN tree (scripts/make_synthetic_srn.py):
what tests/test_gpu_reference_drivers.py and bench.py's reference legs stand on.  CPU only."""
import importlib.util
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_reference_copy_matches_its_manifest_and_the_source_tree():
    mr = _load(os.path.join(ROOT, "oracle", "make_ref.py"), "make_ref")
    assert mr.make(verbose=False) and mr.verify()
    mf = json.load(open(os.path.join(REF, "MANIFEST.json")))
    for rel in ("train.py", "eval.py", "view_synthesis/nerf/__init__.py", "view_synthesis/models/model.py",
                "view_synthesis/utils/util.py", "view_synthesis/datasets/dataset.py", "config/srn-cars-code.yml"):
        assert rel in mf["files"], rel
    src = mf["source"]
    if os.path.isdir(src):                      # build container: byte-identical to the read-only reference
        for rel, meta in mf["files"].items():
            assert mr.sha256(os.path.join(src, rel)) == meta["sha256"], rel


def test_synthetic_srn_tree_loads_through_the_reference_dataset(tmp_path):
    data = _load(os.path.join(ROOT, "scripts", "make_synthetic_srn.py"), "make_srn").main(str(tmp_path), 3, 2, 4, 128)
    code = (
        "import sys, json, numpy as np\n"
        "from view_synthesis.datasets.dataset import SRNDataset\n"
        f"d = SRNDataset({data!r}, 'train'); v = SRNDataset({data!r}, 'val')\n"
        "s = d[5]\n"
        "R, t = s['pose'][:3, :3], s['pose'][:3, 3]\n"
        "print(json.dumps({'n': d.num_objects, 'views': d.num_views, 'nv': v.num_objects, 'shape': list(s['color'].shape),\n"
        "  'K': s['intrinsic'].tolist(), 'orth': float(np.abs(R @ R.T - np.eye(3)).max()), 'rho': float(np.linalg.norm(t)),\n"
        "  'look': float((-R[:, 2] * (-t / np.linalg.norm(t))).sum()), 'fg': float((s['color'][..., :3] < 0.99).any(-1).mean())}))\n")
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([os.path.join(ROOT, "tests", "shims"), REF]), PYTHONDONTWRITEBYTECODE="1")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True).stdout
    r = json.loads(out.strip().splitlines()[-1])
    assert r["n"] == 3 and r["views"] == 4 and r["nv"] == 2
    assert r["shape"] == [96, 96, 4]                               # dataset.py:76-78 crops 1/8 per side
    assert r["K"][0][0] == 131.25 and r["K"][0][2] == 48.0 and r["K"][1][2] == 48.0      # dataset.py:85
    assert r["orth"] < 1e-5 and abs(r["rho"] - 1.3) < 1e-4
    assert r["look"] > 0.999                                       # -z of the camera points at the origin (OpenGL axes)
    assert 0.02 < r["fg"] < 0.9                                    # the object is in view


def test_pythonsafepath_lets_the_drop_in_win_over_the_script_directory(tmp_path):
    """`python <reference>/train.py` puts the reference checkout -- which holds its own `view_synthesis` -- at sys.path[0],
    AHEAD of PYTHONPATH, so the drop-in would never be imported.  PYTHONSAFEPATH=1 (python >= 3.11) removes that entry; the
    recipe of INTEGRATION.md and tests/test_gpu_reference_drivers.py depends on it."""
    ref, drop = tmp_path / "ref", tmp_path / "drop"
    for d, tag in ((ref, "REFERENCE"), (drop, "DROPIN")):
        (d / "view_synthesis").mkdir(parents=True)
        (d / "view_synthesis" / "__init__.py").write_text(f"WHO = {tag!r}\n")
    (ref / "helper.py").write_text("X = 1\n")
    (ref / "train.py").write_text("import view_synthesis, helper\nprint(view_synthesis.WHO)\n")
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([str(drop), str(ref)]))
    env.pop("PYTHONSAFEPATH", None)
    run = lambda e: subprocess.run([sys.executable, str(ref / "train.py")], env=e, capture_output=True, text=True, check=True).stdout.strip()
    assert run(env) == "REFERENCE"
    assert run(dict(env, PYTHONSAFEPATH="1")) == "DROPIN"
