"""bench.py -- rays/s, forward+backward, coarse+fine, of the CodeNeRF volumetric-rendering hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision bf16|fp32]
                    [--stock-caller] [--ddp] [--torch-loss] [--no-prefetch] [--no-graph]

Workload (BASELINE.json configs[1] = SURVEY C2, config/srn-cars-code.yml): per rank and step one loader batch of
B=4 synthetic 128x128 SRN-car-shaped views -> 4 x 4096 random rays = 16,384 rays, processed as the reference's
train.py:92-114 does: 4 chunks of 4096 rays, each chunk = embedding lookup + predict_radiance_and_render (32 coarse +
128 fine samples, hidden 256, codes 256, PE 10/4) + 2 MSE losses + backward + AdamW step.  Ranks are data-parallel:
weak scaling, value = all ranks' rays / max-rank time.

`value`  : inputs (poses, pixel indices, target pixels, object ids) already resident in HBM.
`e2e`    : the same step through the public drop-in API from HOST buffers: pinned images + poses copied H2D, host numpy
           permutation (ray_sampler.sample), target gather, one loss.item() D2H per step -- train.py's loop body, with the
           next step's inputs staged before the current loss is read (--no-prefetch: train.py's order of work).
`--stock-caller`: the loop body an UNCHANGED train.py executes (train.py:96-114, utils/util.py:139-172): two
           torch.nn.functional.mse_loss, loss.item() per CHUNK, the get_params_tensor regulariser (two torch.norm over the
           full tables), torch.optim.AdamW + LambdaLR.step(), DistributedDataParallel x3 when N > 1, no prefetch, no
           codenerf_b200 extras -- only the drop-in view_synthesis.nerf / .models.
`--impl reference`: the UNMODIFIED reference (oracle/_ref, copied by oracle/make_ref.py) through its own
           view_synthesis.nerf / .models API, train.py's loop body, on all host cores (kind="reference"; the oracle port
           only if oracle/_ref is absent).  It runs the C2 shape it prints -- 4 chunks x 4096 rays per step, never a shrunk
           shape -- and caps the NUMBER of timed steps by a time budget (`sample` says how many ran).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "code-nerf_b200")
REF = os.path.join(ROOT, "oracle", "_ref")

CFG = dict(H=128, W=128, B=4, rays_per_image=4096, chunk=4096, Sc=32, Nf=128, near=0.8, far=1.8, hidden=256, code=256,
           Lx=10, Ld=4, num_objects=2458, focal=131.25, lr=1e-4, embedding_lr=1e-3, regularizer_lambda=1e-5)
# algorithmic FLOPs: per sample fwd 558,592; per ray+model fwd 671,232; bwd = 2x fwd - first-layer dX
FLOP_SAMPLE_FWD = 2 * (63 * 256 + 256 * 256 + 256 * 257 + 256 * 256 + 256 * 256 + 256 * 3)
FLOP_SAMPLE_BWD_DATA = FLOP_SAMPLE_FWD - 2 * 63 * 256          # dX of every layer but the first
FLOP_SAMPLE_BWD_W = FLOP_SAMPLE_FWD                            # dW of every layer
FLOP_RAY_FWD = 2 * (3 * 256 * 256 + 256 * 256 + 256 * 257 + 27 * 256 + 3 * 256)
FLOP_PER_RAY_TRAIN = (3 * FLOP_SAMPLE_FWD - 2 * 63 * 256) * (CFG["Sc"] + CFG["Sc"] + CFG["Nf"]) + 2 * 3 * FLOP_RAY_FWD
METRIC = "rays/sec fwd+bwd (coarse+fine)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("CODENERF_PRECISION", "bf16"), choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-eager-baseline", action="store_true", help="skip BASELINE.md row R2 (reference through eager torch on cuda:0)")
    ap.add_argument("--no-hbm-kernels", action="store_true")
    ap.add_argument("--ddp", action="store_true", help="N > 1: wrap the modules in DistributedDataParallel as "
                    "utils/util.py:139-142 does, instead of codenerf_b200.parallel.GradAllReduce")
    ap.add_argument("--torch-loss", action="store_true", help="2 x torch.nn.functional.mse_loss + add (train.py:103-104 verbatim) "
                    "instead of codenerf_b200.ops.mse_pair")
    ap.add_argument("--staged-loss", action="store_true", help="codenerf_b200.ops.mse_pair on the returned rgb tensors instead of "
                    "the loss folded into the compositing epilogues (nerf.predict_radiance_render_and_loss)")
    ap.add_argument("--no-graph", action="store_true", help="enqueue every step from Python instead of replaying a CUDA graph of it")
    ap.add_argument("--no-prefetch", action="store_true", help="e2e leg: draw each step's pixel permutations at the start "
                    "of the step (the reference's order of work) instead of under the previous step's kernels")
    ap.add_argument("--stock-caller", action="store_true", help="train.py:96-114 verbatim around the drop-in (see module docstring)")
    # reference arm
    ap.add_argument("--ref-device", default="cpu", help="--impl reference: cpu (R1, the arm the driver runs) or cuda:0 (R2)")
    ap.add_argument("--ref-budget-s", type=float, default=170.0, help="--impl reference: wall-clock budget for the timed steps")
    return ap.parse_args()


def look_at_pose(theta, phi, rho):
    """camera-to-world on a sphere looking at the origin (closed form of eval.py:22-38)."""
    c2w = np.eye(4, dtype=np.float32)
    c2w[0, 0], c2w[1, 0] = -np.sin(phi), np.cos(phi)
    c2w[0, 1], c2w[1, 1], c2w[2, 1] = -np.sin(theta) * np.cos(phi), -np.sin(theta) * np.sin(phi), np.cos(theta)
    c2w[0, 2], c2w[1, 2], c2w[2, 2] = np.cos(theta) * np.cos(phi), np.cos(theta) * np.sin(phi), np.sin(theta)
    c2w[:3, 3] = rho * c2w[:3, 2]
    return c2w


def synthetic_batch(rng, c=CFG):
    """One loader batch of SRN-car-shaped data: images (B,H,W,4) in [0,1], poses (B,4,4), object ids (B,)."""
    B, H, W = c["B"], c["H"], c["W"]
    poses = np.stack([look_at_pose(rng.uniform(0.2, 1.4), rng.uniform(-np.pi, np.pi), 1.3) for _ in range(B)])
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    imgs = np.ones((B, H, W, 4), dtype=np.float32)
    for b in range(B):
        cx, cy, r = rng.uniform(48, 80), rng.uniform(48, 80), rng.uniform(20, 36)
        inside = ((xx - cx) ** 2 / r ** 2 + (yy - cy) ** 2 / (0.6 * r) ** 2) < 1.0
        shade = 0.5 + 0.5 * np.cos((xx + yy) / 17.0 + b)
        for ch in range(3):
            imgs[b, ..., ch] = np.where(inside, shade * rng.uniform(0.2, 1.0), 1.0)
    ids = rng.integers(0, c["num_objects"], size=(B,)).astype(np.int64)
    return imgs, poses, ids


def intrinsics(c=CFG):
    return torch.tensor([[c["focal"], 0, c["W"] / 2, 0], [0, c["focal"], c["H"] / 2, 0], [0, 0, 1, 0], [0, 0, 0, 1]],
                        dtype=torch.float32)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=lambda: [self.rows.append(l) for l in self.proc.stdout], daemon=True).start()
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for l in self.rows:
            f = [x.strip() for x in l.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def workload_config(precision):
    c = CFG
    return {"workload": "srn-cars-code training (SURVEY C2 / BASELINE configs[1]): synthetic 128x128 SRN-car views, "
                        "B=4 images x 4096 rays = 16384 rays per rank per step in 4 chunks of 4096, 32 coarse + 128 fine "
                        "samples, CodeNeRF hidden 256 / codes 256, PE L=10/4, per chunk: embedding lookup + render fwd + "
                        "2 MSE + backward + AdamW step",
            "rays_per_step_per_gpu": c["B"] * c["rays_per_image"], "chunk": c["chunk"], "mlp_precision": precision,
            "pairing": "reference"}


# =====================================================================================================================
# Reference arm: the unmodified reference through its own API (oracle/_ref), train.py's loop body
# =====================================================================================================================
def load_reference_modules():
    """-> (nerf, models, kind).  The live reference from oracle/_ref with `imageio` stubbed (only the dataset module, which
    this arm does not use, imports it -- SURVEY 8c); the oracle port only when the copy is absent."""
    if os.path.isdir(os.path.join(REF, "view_synthesis")):
        sys.dont_write_bytecode = True
        sys.modules.setdefault("imageio", types.ModuleType("imageio"))
        sys.path.insert(0, REF)
        import view_synthesis.nerf as nerf
        import view_synthesis.models as models
        assert os.path.realpath(nerf.__file__).startswith(os.path.realpath(REF)), nerf.__file__
        return nerf, models, "reference"
    return None, None, "port"


def reference_trainer(device, c=CFG, seed=56):
    """Builds the reference objects as train.py:42-58 / utils/util.py:93-172 do and returns step(imgs, poses, ids) -> the
    loop body of train.py:70-114 for one loader batch (4 chunk steps).  Falls back to the oracle port when the reference
    copy is missing."""
    nerf, models, kind = load_reference_modules()
    np.random.seed(seed); torch.manual_seed(seed)
    K = intrinsics(c)
    n_img, chunk = c["rays_per_image"], c["chunk"]
    if kind == "reference":
        emb = models.ShapeTextureEmbedding(c["num_objects"], c["code"], c["code"]).to(device)
        mk = lambda: models.CodeNeRFModel(hidden_size=c["hidden"], num_embeddings=c["num_objects"], shape_code_size=c["code"],
                                          texture_code_size=c["code"], num_encoding_fn_xyz=c["Lx"], include_input_xyz=True,
                                          num_encoding_fn_dir=c["Ld"], include_input_dir=True).to(device)
        coarse, fine = mk(), mk()
        ray_sampler = nerf.RaySampler(c["H"], c["W"], K, sample_size=n_img, device=device, datatype=torch.float32)
        point_sampler = nerf.PointSampler(c["Sc"], c["Nf"], c["near"], c["far"], spacing_mode="lindepth", perturb=True,
                                          dtype=torch.float32, device=device)
        embedders = (nerf.PositionalEmbedder(num_freq=c["Lx"], log_sampling=True, include_input=True, dtype=torch.float32, device=device),
                     nerf.PositionalEmbedder(num_freq=c["Ld"], log_sampling=True, include_input=True, dtype=torch.float32, device=device))
        opt = torch.optim.AdamW([{"params": coarse.parameters()}, {"params": fine.parameters()},
                                 {"params": emb.parameters(), "lr": c["embedding_lr"]}], lr=c["lr"])          # utils/util.py:159-164
        sched = torch.optim.lr_scheduler.LambdaLR(opt, lr_lambda=lambda e: 0.1 ** (e / 5000000))               # :166-170

        def step(imgs, poses, ids):
            ro_b, rd_b, sel = ray_sampler.sample(tform_cam2world=poses)                                         # train.py:76
            tgt = imgs.flatten(1, 2)
            oid = ids[:, None].expand(-1, n_img)
            tgt = torch.cat([tgt[k, sel[k], :] for k in range(c["B"])], dim=0)
            oid = oid.reshape(-1)
            last = 0.0
            for i in range(0, ro_b.shape[0], chunk):                                                            # train.py:92-114
                z = emb(oid[i:i + chunk])
                rc, rf = nerf.predict_radiance_and_render((ro_b[i:i + chunk], rd_b[i:i + chunk]), point_sampler, embedders,
                                                          coarse, fine, z)
                lc = torch.nn.functional.mse_loss(rc[..., :3], tgt[i:i + chunk, :3])
                lf = torch.nn.functional.mse_loss(rf[..., :3], tgt[i:i + chunk, :3])
                last = lf.item()                                                                                # :105
                sp, tp = models.get_params_tensor(emb, False)
                loss = lc + lf + c["regularizer_lambda"] * (torch.norm(sp, p=2) + torch.norm(tp, p=2))
                opt.zero_grad()
                loss.backward()
                opt.step()
                sched.step()
            return last
        return step, kind
    # ---- oracle port (no reference copy on this machine)
    sys.path.insert(0, ROOT)
    from oracle import codenerf_oracle as orc
    cfg = orc.RenderConfig(c["Sc"], c["Nf"], c["near"], c["far"], "lindepth", c["Lx"], c["Ld"])
    for k in ("freqs_xyz", "freqs_dir", "z", "lower", "upper"):
        setattr(cfg, k, getattr(cfg, k).to(device))
    pc = {k: v.to(device).requires_grad_(True) for k, v in orc.make_codenerf_params(1, c["hidden"], c["code"]).items()}
    pf = {k: v.to(device).requires_grad_(True) for k, v in orc.make_codenerf_params(2, c["hidden"], c["code"]).items()}
    tab_s = torch.randn(c["num_objects"], c["code"], device=device).requires_grad_(True)
    tab_t = torch.randn(c["num_objects"], c["code"], device=device).requires_grad_(True)
    opt = torch.optim.AdamW([{"params": list(pc.values())}, {"params": list(pf.values())},
                             {"params": [tab_s, tab_t], "lr": c["embedding_lr"]}], lr=c["lr"])
    d = orc.camera_directions(c["H"], c["W"], K).to(device)

    def step(imgs, poses, ids):
        sel = orc.draw_select_inds(c["H"] * c["W"], c["B"], n_img)
        ro_b, rd_b = orc.sample_rays(d, poses, sel)
        tgt = torch.cat([imgs.flatten(1, 2)[b, sel[b], :3] for b in range(c["B"])])
        oid = ids[:, None].expand(-1, n_img).reshape(-1)
        last = 0.0
        for i in range(0, ro_b.shape[0], chunk):
            z_s, z_t = orc.embedding_lookup(tab_s, tab_t, oid[i:i + chunk])
            n = min(chunk, ro_b.shape[0] - i)
            rc, rf, _ = orc.predict_radiance_and_render(cfg, ro_b[i:i + chunk], rd_b[i:i + chunk], pc, pf, z_s, z_t,
                                                        torch.rand(n, c["Sc"], device=device), torch.rand(n, c["Nf"], device=device))
            lc, lf = torch.nn.functional.mse_loss(rc, tgt[i:i + chunk]), torch.nn.functional.mse_loss(rf, tgt[i:i + chunk])
            last = lf.item()
            loss = lc + lf + c["regularizer_lambda"] * (torch.norm(tab_s.detach()) + torch.norm(tab_t.detach()))
            opt.zero_grad()
            loss.backward()
            opt.step()
        return last
    return step, kind


def time_reference(device, steps, warmup, budget_s, c=CFG):
    """-> dict(rays_per_s, sec_per_step, steps_run, warmup_run, threads, kind).  Every step is the full C2 step (16,384 rays in
    4 chunks of 4096); only the NUMBER of timed steps is capped by the budget (at least 2 on the CPU)."""
    device = torch.device(device)
    threads = os.cpu_count() or 1
    if device.type == "cpu":
        torch.set_num_threads(threads)
    step, kind = reference_trainer(device, c)
    rng = np.random.default_rng(0)
    imgs, poses, ids = (torch.from_numpy(a).to(device) for a in synthetic_batch(rng, c))
    sync = (lambda: torch.cuda.synchronize(device)) if device.type == "cuda" else (lambda: None)
    warm = max(1, min(warmup, 1 if device.type == "cpu" else 3))
    t0 = time.perf_counter()
    for _ in range(warm):
        step(imgs, poses, ids)
    sync()
    est = (time.perf_counter() - t0) / warm
    floor = 2 if device.type == "cpu" else 10
    n = int(max(floor, min(steps, budget_s // max(est, 1e-6)))) if steps >= floor else max(1, steps)
    t0 = time.perf_counter()
    for _ in range(n):
        step(imgs, poses, ids)
    sync()
    sec = (time.perf_counter() - t0) / n
    rays = c["B"] * c["rays_per_image"]
    return {"rays_per_s": rays / sec, "sec_per_step": sec, "steps_run": n, "warmup_run": warm, "threads": threads, "kind": kind}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = time_reference(args.ref_device, args.steps, args.warmup, args.ref_budget_s)
    on_cpu = torch.device(args.ref_device).type == "cpu"
    where = f"{r['threads']} host threads (torch CPU, fp32)" if on_cpu else f"stock eager torch on {args.ref_device} (fp32, TF32 off)"
    src = ("the unmodified reference (oracle/_ref: view_synthesis.nerf / .models), train.py:70-114 loop body" if r["kind"] == "reference"
           else "oracle port of the reference path (oracle/_ref absent)")
    sample = (f"{r['steps_run']} timed steps (of {args.steps} requested; capped by a {args.ref_budget_s:.0f} s budget) after {r['warmup_run']} "
              f"warm-up, each the FULL C2 step: 16384 rays = 4 chunks x 4096 (32+128 samples): sample + gather + 4 x (embedding + render "
              f"fwd + 2 MSE + .item() + regulariser + backward + AdamW + LambdaLR); {src}; {where}")
    line = {"impl": "reference", "metric": METRIC, "value": r["rays_per_s"], "unit": "rays/s",
            "n_gpus": args.gpus, "steps": r["steps_run"], "warmup": r["warmup_run"], "ms_per_step": r["sec_per_step"] * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config("fp32"), device=args.ref_device,
                           note="N > 1: rank 0 alone runs this arm -- the driver's N-GPU ratio is N GPUs over ONE CPU process"),
            "gpu_launches": 0,
            "cpu_baseline": {"value": r["rays_per_s"], "unit": "rays/s", "cores": r["threads"] if on_cpu else 0, "kind": r["kind"],
                             "sample": sample},
            "e2e": {"value": r["rays_per_s"], "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def reference_subprocess(device, steps, budget_s, timeout=900):
    """Runs `bench.py --impl reference` in a child process (the reference's `view_synthesis` package cannot share a process
    with the drop-in of the same name) and returns its parsed JSON line, or {"error": ...}."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--ref-device", device, "--steps", str(steps),
           "--warmup", "1", "--ref-budget-s", str(budget_s)]
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
        for l in reversed(out.stdout.strip().splitlines()):
            if l.startswith("{"):
                return json.loads(l)
        return {"error": (out.stderr or out.stdout)[-400:]}
    except Exception as e:      # noqa: BLE001
        return {"error": repr(e)}


# =====================================================================================================================
def run_ours(args):
    sys.path.insert(0, PKG)
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import view_synthesis.nerf as nerf
    import view_synthesis.models as models
    from codenerf_b200 import _lib, ops
    c = CFG
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device (no CPU fallback)")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    _lib.check(_lib.load().cnrf_device_check(), "cnrf_device_check")
    nerf.set_precision(args.precision)
    seed = 55 + rank + 1                                                  # train.py:29-31
    np.random.seed(seed); torch.manual_seed(seed)
    rng = np.random.default_rng(seed)
    stock = args.stock_caller
    # CUDA graph of the whole step (4 chunks: embedding lookup, render + loss, backward, gradient exchange, AdamW): the extras
    # caller only (the stock loop reads a loss on the host in every chunk); torch.cuda.graph around the same Python code
    use_graph = not (stock or args.no_graph or args.ddp or args.torch_loss)
    use_ddp = world > 1 and (args.ddp or stock)

    mods = {"embedding": models.ShapeTextureEmbedding(c["num_objects"], c["code"], c["code"]).to(dev),
            "nerf_coarse": models.CodeNeRFModel(c["hidden"], c["num_objects"], c["code"], c["code"], c["Lx"], c["Ld"]).to(dev),
            "nerf_fine": models.CodeNeRFModel(c["hidden"], c["num_objects"], c["code"], c["code"], c["Lx"], c["Ld"]).to(dev)}
    sync = None
    if use_ddp:                                                           # utils/util.py:139-142, as the reference wraps them
        from torch.nn.parallel import DistributedDataParallel as DDP
        mods = {k: DDP(m, device_ids=[local], output_device=local) for k, m in mods.items()}
    elif world > 1:
        # the same exchange (mean of every MLP-weight and code-table gradient over ranks, rank 0's initial parameters) as ONE
        # flat NCCL all-reduce per chunk step instead of three DDP reducers: codenerf_b200/parallel.py
        from codenerf_b200.parallel import GradAllReduce
        sync = GradAllReduce([p for m in mods.values() for p in m.parameters()])
    groups = [{"params": mods["nerf_coarse"].parameters()}, {"params": mods["nerf_fine"].parameters()},
              {"params": mods["embedding"].parameters(), "lr": c["embedding_lr"]}]
    if stock:
        opt = torch.optim.AdamW(groups, lr=c["lr"])                                                          # utils/util.py:159-164
        sched = torch.optim.lr_scheduler.LambdaLR(opt, lr_lambda=lambda e: 0.1 ** (e / 5000000))            # :166-170
    else:
        from codenerf_b200.optim import FusedAdamW          # torch.optim.AdamW arithmetic, one launch
        opt, sched = FusedAdamW(groups, lr=c["lr"], capturable=use_graph), None
    K = intrinsics(c)
    ray_sampler = nerf.RaySampler(c["H"], c["W"], K, c["rays_per_image"], dev, torch.float32)
    point_sampler = nerf.PointSampler(c["Sc"], c["Nf"], c["near"], c["far"], "lindepth", True, torch.float32, dev)
    embedders = (nerf.PositionalEmbedder(c["Lx"], True, True, torch.float32, dev),
                 nerf.PositionalEmbedder(c["Ld"], True, True, torch.float32, dev))
    imgs_np, poses_np, ids_np = synthetic_batch(rng, c)
    imgs_pin = torch.from_numpy(imgs_np).pin_memory()
    poses_pin = torch.from_numpy(poses_np).pin_memory()
    ids_pin = torch.from_numpy(ids_np).pin_memory()
    n_rays = c["B"] * c["rays_per_image"]
    fused_loss = not (stock or args.torch_loss or args.staged_loss)

    def chunks_step(ro, rd, tgt, oid, read_loss):
        total = None
        for i in range(0, n_rays, c["chunk"]):
            sl = slice(i, i + c["chunk"])
            codes = mods["embedding"](oid[sl])
            if fused_loss:       # the loss rides in the compositing epilogues (cnrf_render_loss_fwd / _bwd)
                lc, lf, loss, _rc, _rf = nerf.predict_radiance_render_and_loss(
                    (ro[sl], rd[sl]), point_sampler, embedders, mods["nerf_coarse"], mods["nerf_fine"], codes, target_pixels=tgt[sl])
                opt.zero_grad()
                loss.backward()
                if sync is not None:
                    sync()
                opt.step()
                if read_loss:
                    total = lf.detach() if total is None else total + lf.detach()
                continue
            rgb_c, rgb_f = nerf.predict_radiance_and_render((ro[sl], rd[sl]), point_sampler, embedders,
                                                            mods["nerf_coarse"], mods["nerf_fine"], codes)
            if stock:                                                    # train.py:103-114 verbatim
                lc = torch.nn.functional.mse_loss(rgb_c[..., :3], tgt[sl][..., :3])
                lf = torch.nn.functional.mse_loss(rgb_f[..., :3], tgt[sl][..., :3])
                _psnr = -10.0 * np.log10(max(lf.item(), 1e-12))          # utils.mse2psnr(nerf_loss_fine.item()): a sync per CHUNK
                sp, tp = models.get_params_tensor(mods["embedding"], use_ddp)
                loss = lc + lf + c["regularizer_lambda"] * (torch.norm(sp, p=2) + torch.norm(tp, p=2))
                opt.zero_grad()
                loss.backward()
                opt.step()
                sched.step()
                continue
            if args.torch_loss:
                lc = torch.nn.functional.mse_loss(rgb_c[..., :3], tgt[sl])
                lf = torch.nn.functional.mse_loss(rgb_f[..., :3], tgt[sl])
                loss = lc + lf
            else:                                                        # the same three numbers, one launch
                lc, lf, loss = ops.mse_pair(rgb_c, rgb_f, tgt[sl])
            opt.zero_grad()
            loss.backward()
            if sync is not None:
                sync()
            opt.step()
            if read_loss:
                total = lf.detach() if total is None else total + lf.detach()
        return total

    # device-resident inputs for `value`
    imgs_d, poses_d, ids_d = imgs_pin.to(dev), poses_pin.to(dev), ids_pin.to(dev)
    sel_np = np.stack([np.random.permutation(c["H"] * c["W"])[:c["rays_per_image"]] for _ in range(c["B"])])
    sel_d = torch.from_numpy(sel_np).to(dev)
    tgt_d = torch.cat([imgs_d.flatten(1, 2)[b, sel_d[b], :3] for b in range(c["B"])]).contiguous()
    oid_d = ids_d[:, None].expand(-1, c["rays_per_image"]).reshape(-1)

    def step_resident():
        ro, rd = ops.ray_bundle(poses_d, c["H"], c["W"], c["focal"], c["W"] / 2, c["H"] / 2, select=sel_d,
                                n=c["rays_per_image"])
        chunks_step(ro, rd, tgt_d, oid_d, False)

    h2d = imgs_pin.numel() * 4 + poses_pin.numel() * 4 + ids_pin.numel() * 8 + n_rays * 8
    prefetch = not (args.no_prefetch or stock)
    staged = []

    def stage_inputs():
        """One step's host -> device work: pinned images / poses / ids H2D, ray_sampler.sample (host numpy permutation +
        ray kernel), target-pixel gather (train.py:76-80)."""
        imgs = imgs_pin.to(dev, non_blocking=True)
        poses = poses_pin.to(dev, non_blocking=True)
        ids = ids_pin.to(dev, non_blocking=True)
        ro, rd, _sel = ray_sampler.sample(tform_cam2world=poses)         # train.py:76
        sel = _sel if stock else ray_sampler.select_inds_device          # stock caller: the numpy indices, as train.py:79 uses them
        flat = imgs.flatten(1, 2)
        tgt = torch.cat([flat[k, sel[k], :] for k in range(c["B"])], dim=0)   # train.py:77-80
        if not stock:
            tgt = tgt[..., :3]
        oid = ids[:, None].expand(-1, c["rays_per_image"]).reshape(-1)
        staged.append((ro, rd, tgt, oid))

    def step_e2e():
        # Every step does one stage_inputs() and one loss read.  With prefetching (default) the inputs of step k+1 are staged
        # after step k's kernels have been enqueued and before its loss is read -- a prefetching loader -- so the host
        # permutation and the copies run under the GPU's work; --no-prefetch keeps train.py's order (stage, compute, read).
        if not prefetch or not staged:
            stage_inputs()
        total = chunks_step(*staged.pop(0), True)
        if prefetch:
            stage_inputs()
        if total is not None:
            _ = total.item()               # the step's result: summed fine loss of its chunks, one device->host read per step

    graphs = {}

    def build_graphs():
        """Capture (a) the device-resident step and (b) the chunk loop of the end-to-end step over static input buffers.
        Called after eager warm-up steps (optimiser state, per-device attributes and NCCL are initialised outside capture)."""
        st_in = (torch.empty(n_rays, 3, device=dev), torch.empty(n_rays, 3, device=dev), torch.empty(n_rays, 3, device=dev),
                 torch.empty(n_rays, dtype=torch.int64, device=dev))
        torch.cuda.synchronize()
        g_res = torch.cuda.CUDAGraph()
        opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(g_res):
            step_resident()
        g_e2e = torch.cuda.CUDAGraph()
        opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(g_e2e):
            total = chunks_step(*st_in, True)
        torch.cuda.synchronize()
        graphs.update(resident=g_res, e2e=g_e2e, inputs=st_in, total=total)

    def step_resident_graph():
        graphs["resident"].replay()

    def step_e2e_graph():
        # the same order of work as step_e2e; the staged inputs are copied into the graph's static buffers (device to device)
        if not prefetch or not staged:
            stage_inputs()
        for dst, src in zip(graphs["inputs"], staged.pop(0)):
            dst.copy_(src, non_blocking=True)
        graphs["e2e"].replay()
        if prefetch:
            stage_inputs()
        _ = graphs["total"].item()         # the step's result: summed fine loss of its chunks, one device->host read per step

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # standalone rooflines of the HBM-bound stages (north_star: compositing / sampling vs the HBM peak): one launch per
    # repetition over 524,288 rays (working sets > L2), CUDA events, algorithmic bytes of SURVEY section 8(d).  Measured BEFORE
    # the training steps: these kernels are issue-bound, and after a minute of graph-replayed tensor work the box sits at its
    # power cap (SM clock ~1.82 GHz instead of 1.965), which costs them 7-8 % that have nothing to do with the kernels.
    hbm_rows, hbm_peak = None, None
    if world == 1 and rank == 0 and not args.no_hbm_kernels:
        import importlib.util
        spec = importlib.util.spec_from_file_location("time_hbm_kernels", os.path.join(ROOT, "scripts", "time_hbm_kernels.py"))
        thk = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(thk)
        hbm_rows = [{k: (round(r[k], 4) if isinstance(r[k], float) else r[k]) for k in
                     ("kernel", "rays", "bytes_per_ray", "ms", "GBps", "frac")}
                    for r in thk.measure(524288, 10, dev, verbose=False)]
        hbm_peak = {"GBps": thk.hbm_peak()[0], "source": thk.hbm_peak()[1]}
        torch.cuda.empty_cache()
    lib = _lib.load()
    step_resident(); torch.cuda.synchronize()
    n0 = lib.cnrf_launch_count(); step_resident(); torch.cuda.synchronize()
    launches_per_step = int(lib.cnrf_launch_count() - n0)
    clocks = ClockSampler(local)
    clocks.start()
    fine_rows = c["chunk"] * (c["Sc"] + c["Nf"])
    tc = args.precision == "bf16"
    if tc:
        for which in range(3):
            lib.cnrf_debug_kernel_ms(which, fine_rows, None, None)       # time the fine-pass launches of the three chain kernels
    graph_note = None
    if use_graph:
        try:
            for which in range(3):
                lib.cnrf_debug_kernel_ms(which, 0, None, None)           # no events while capturing
            build_graphs()
        except Exception as e:      # noqa: BLE001  (a capture that fails is reported, and the step is enqueued from Python)
            graphs.clear()
            graph_note = "capture failed (%s): eager launches" % repr(e)[:200]
            torch.cuda.synchronize()
        if tc:
            for which in range(3):
                lib.cnrf_debug_kernel_ms(which, fine_rows, None, None)
    if graphs:
        # timed region: graph replays.  CUDA events cannot be recorded inside a captured graph, so the per-kernel durations
        # of the roofline come from an eager pass of the same steps run right after, in the same process and clock window.
        for which in range(3):
            lib.cnrf_debug_kernel_ms(which, 0, None, None)
        ms_val = timed(step_resident_graph, args.steps, max(3, args.warmup))
        if tc:
            for which in range(3):
                lib.cnrf_debug_kernel_ms(which, fine_rows, None, None)
        ms_eager = timed(step_resident, args.steps, 1)
        kernel_ms = [float(lib.cnrf_debug_kernel_ms(w, 0, None, None)) if tc else 0.0 for w in range(3)]
        clk = clocks.stop()
        ms_e2e = timed(step_e2e_graph, args.steps, 2)
    else:
        ms_val = timed(step_resident, args.steps, max(3, args.warmup))
        ms_eager = ms_val
        kernel_ms = [float(lib.cnrf_debug_kernel_ms(w, 0, None, None)) if tc else 0.0 for w in range(3)]
        clk = clocks.stop()
        ms_e2e = timed(step_e2e, args.steps, 1)
    d2h = 4 * (n_rays // c["chunk"] if stock else 1)
    total_rays = n_rays * world * args.steps
    value, e2e = total_rays / (ms_val * 1e-3), total_rays / (ms_e2e * 1e-3)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        roof, others, chunk_traffic = kernel_rooflines(args, kernel_ms, ms_eager / args.steps, clk, peaks)
        cfgd = workload_config(args.precision)
        cfgd.update({
            "caller": ("stock train.py loop body (train.py:96-114): 2 x mse_loss, .item() per chunk, get_params_tensor regulariser, "
                       "torch.optim.AdamW + LambdaLR, no prefetch" if stock else
                       "codenerf_b200 extras: " + ("loss folded into the compositing epilogues (predict_radiance_render_and_loss)" if not (args.torch_loss or args.staged_loss) else "ops.mse_pair" if not args.torch_loss else "2 x torch mse_loss (--torch-loss)") +
                       ", FusedAdamW (one launch), one loss read per step"),
            "gradient_exchange": ("none (1 GPU)" if world == 1 else "DistributedDataParallel x3 (utils/util.py:139-142)" if use_ddp
                                  else "codenerf_b200.parallel.GradAllReduce (one flat NCCL all-reduce per chunk step)"),
            "optimizer": "AdamW lr 1e-4, embeddings 1e-3 (train.py:111-114): " + ("torch.optim.AdamW" if stock else "codenerf_b200.optim.FusedAdamW"),
            "e2e_host_overlap": ("prefetching loader: the next step's inputs (pinned H2D copies, np.random.permutation draws from the same RNG "
                                 "stream, ray kernel, target gather) are staged after this step's kernels are enqueued and before its loss "
                                 "is read" if prefetch else "off: train.py's order of work (stage, compute, read)"),
            "l2": "activations written/read per chunk (>1 GB) exceed the 126 MB L2; no explicit flush",
            "launch_mode": ("CUDA graph of the whole step (4 chunks), replayed once per step; FusedAdamW(capturable=True) keeps its step "
                            "count on the device; the roofline's per-kernel durations come from an eager pass of the same steps "
                            "(%.3f ms per step) run right after the timed region" % (ms_eager / args.steps)) if graphs else
                           (graph_note or "eager: every launch enqueued from Python")})
        line = {"metric": METRIC, "value": value, "unit": "rays/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_val / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "bf16" if tc else "f32", "data": "synthetic", "config": cfgd, "clocks": clk,
                "e2e": {"value": e2e, "unit": "rays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "gpu_launches": launches_per_step, "roofline": roof, "roofline_others": others,
                "chain_kernels_dram_bytes_per_chunk": chunk_traffic,
                "model_flops_per_ray": FLOP_PER_RAY_TRAIN,
                "achieved_tflops_whole_step": value * FLOP_PER_RAY_TRAIN / 1e12 / world}
        if hbm_rows is not None:
            line["hbm_kernels"] = hbm_rows
            line["hbm_kernels_peak"] = hbm_peak
        torch.cuda.empty_cache()
        if world == 1 and not args.no_eager_baseline:
            # the unmodified reference through stock eager torch on this B200, >= 10 full C2 steps
            r2 = reference_subprocess(f"cuda:{local}", 10, 60.0)
            line["eager_torch_b200"] = ({"value": r2["value"], "unit": "rays/s", "steps": r2["steps"], "ms_per_step": r2["ms_per_step"],
                                         "kind": r2["cpu_baseline"]["kind"], "what": "BASELINE.md R2: " + r2["cpu_baseline"]["sample"]}
                                        if "value" in r2 else r2)
        if world == 1 and not args.no_cpu_baseline:
            # one full C2 step (4 chunks x 4096 rays) after one warm-up step, on all host cores
            r1 = reference_subprocess("cpu", 1, 1.0)
            line["cpu_baseline"] = ({"value": r1["value"], "unit": "rays/s", "cores": r1["cpu_baseline"]["cores"],
                                     "kind": r1["cpu_baseline"]["kind"], "sample": r1["cpu_baseline"]["sample"]}
                                    if "value" in r1 else {"value": None, "unit": "rays/s", "cores": os.cpu_count(), "kind": "reference",
                                                           "sample": "failed: " + r1.get("error", "")})
        print(json.dumps(line), flush=True)
    if world > 1:
        if graphs:
            # NCCL all-reduces live inside the captured graphs: tearing the communicator down with them alive hung at
            # interpreter exit when this was first measured.  Release the graphs, drain the device, agree that every rank
            # is done, and leave without running the remaining destructors.
            graphs.clear()
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            sys.stdout.flush(); sys.stderr.flush()
            os._exit(0)
        dist.destroy_process_group()


def _ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per fine-pass launch of the three chain kernels, from the committed
    `ncu --set full` capture of this round (profiles/r02_tc_kernels_ncu.json), else last round's forward-only figure."""
    for name in ("r02_tc_kernels_ncu.json",):
        try:
            return json.load(open(os.path.join(ROOT, "profiles", name)))
        except Exception:
            pass
    try:
        return {"mlp_fwd_tc_kernel": {"dram_bytes_per_launch": json.load(open(os.path.join(ROOT, "profiles", "r01_fwd_kernel_ncu.json")))["dram_bytes_per_launch"]}}
    except Exception:
        return {}


def kernel_rooflines(args, kernel_ms, ms_per_step, clk, peaks, c=CFG):
    """Times of the three tensor-core chain kernels on the fine pass (4096 rays x 160 samples), measured live by CUDA events
    on the launching stream around every such launch of the timed region (cnrf_debug_kernel_ms).  `roofline` = the kernel
    with the largest share of the step; the others go to `roofline_others`.  Denominator: the measured bf16 burst peak when
    the sampled SM clock of the timed window is at the burst clock (>= 1.9 GHz), the sustained peak otherwise."""
    N, S = c["chunk"], c["Sc"] + c["Nf"]
    rows = N * S
    chunks = c["B"] * c["rays_per_image"] // c["chunk"]
    sm = (clk or {}).get("sm_mhz") or 0.0
    burst = sm >= 1900.0 and "bf16_tflops" in peaks
    key = "bf16_tflops" if burst else ("bf16_tflops_sustained" if "bf16_tflops_sustained" in peaks else "bf16_tflops")
    peak = peaks.get(key, 1590.0 if burst else 1400.0)
    peak_src = (f"MEASURED_PEAKS.json {key}" if key in peaks else "fallback (B200_PROFILING.md)") + \
        f" -- chosen from the sampled SM clock of the timed window ({sm:.0f} MHz)"
    traffic = _ncu_traffic()
    names = ("mlp_fwd_tc_kernel", "mlp_bwd_tc_kernel", "mlp_dw_tc_kernel")
    flops = (rows * FLOP_SAMPLE_FWD, rows * FLOP_SAMPLE_BWD_DATA, rows * FLOP_SAMPLE_BWD_W)
    what = ("forward chain", "backward-data chain", "weight gradients")
    out = []
    for i in range(3):
        ms = kernel_ms[i]
        ach = flops[i] / (ms * 1e-3) / 1e12 if ms > 0 else 0.0
        tr = (traffic.get(names[i]) or {}).get("dram_bytes_per_launch")
        d = {"bound": "tensor", "kernel": f"{names[i]} ({what[i]}), fine pass: {N} rays x {S} samples = {rows} sample rows",
             "achieved": ach, "peak": peak, "peak_source": peak_src, "unit": "TFLOP/s", "frac": ach / peak if peak else 0.0,
             "traffic": tr, "ms_per_launch": ms, "algorithmic_flops_per_launch": flops[i],
             "share_of_step": (ms * chunks / ms_per_step) if ms_per_step > 0 else 0.0}
        if tr and ms > 0:
            d["dram_GBps"] = tr / (ms * 1e-3) / 1e9
            d["dram_frac_of_hbm_peak"] = d["dram_GBps"] / peaks.get("hbm_gbs", 6650.0)
        out.append(d)
    if args.precision != "bf16":
        return ({"bound": "tensor", "kernel": "fp32 SIMT chain (parity mode)", "achieved": 0.0, "peak": peak, "peak_source": peak_src,
                 "unit": "TFLOP/s", "frac": 0.0, "traffic": None}, [], None)
    order = sorted(range(3), key=lambda i: -out[i]["share_of_step"])
    chunk_traffic = None
    if all((traffic.get(n) or {}).get("dram_bytes_per_launch") for n in names):
        # fine-pass launch bytes scaled to fine + coarse rows (the kernels' traffic is proportional to sample rows)
        chunk_traffic = sum(traffic[n]["dram_bytes_per_launch"] for n in names) * (S + c["Sc"]) / S
    return out[order[0]], [out[i] for i in order[1:]], chunk_traffic


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    run_ours(args)


if __name__ == "__main__":
    main()
