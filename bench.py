"""bench.py -- rays/s, forward+backward, coarse+fine, of the CodeNeRF volumetric-rendering hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision bf16|fp32]

Workload (BASELINE.json configs[1] = SURVEY C2, config/srn-cars-code.yml): per rank and step one loader batch of
B=4 synthetic 128x128 SRN-car-shaped views -> 4 x 4096 random rays = 16,384 rays, processed as the reference's
train.py:92-114 does: 4 chunks of 4096 rays, each chunk = embedding lookup + predict_radiance_and_render (32 coarse +
128 fine samples, hidden 256, codes 256, PE 10/4) + 2 MSE losses + backward + AdamW step.  Ranks are data-parallel
(DistributedDataParallel over NCCL, as utils/util.py:139-142): weak scaling, value = all ranks' rays / max-rank time.

`value`  : inputs (poses, pixel indices, target pixels, object ids) already resident in HBM.
`e2e`    : the same step through the public drop-in API from HOST buffers: pinned images + poses copied H2D, host numpy
           permutation (ray_sampler.sample), target gather, one loss.item() D2H per step -- train.py's loop body, with the
           next step's inputs staged before the current loss is read (--no-prefetch: train.py's order of work).
`--impl reference`: the CPU oracle restatement of the reference path (the reference itself is Python and cannot travel to
           the GPU box; kind="port") on all host cores, bounded sample per step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200"))
sys.path.insert(0, ROOT)

CFG = dict(H=128, W=128, B=4, rays_per_image=4096, chunk=4096, Sc=32, Nf=128, near=0.8, far=1.8, hidden=256, code=256,
           Lx=10, Ld=4, num_objects=2458, focal=131.25)
# algorithmic FLOPs: per sample fwd 558,592; per ray+model fwd 671,232; bwd = 2x fwd - first-layer dX
FLOP_SAMPLE_FWD = 2 * (63 * 256 + 256 * 256 + 256 * 257 + 256 * 256 + 256 * 256 + 256 * 3)
FLOP_RAY_FWD = 2 * (3 * 256 * 256 + 256 * 256 + 256 * 257 + 27 * 256 + 3 * 256)
FLOP_PER_RAY_TRAIN = (3 * FLOP_SAMPLE_FWD - 2 * 63 * 256) * (CFG["Sc"] + CFG["Sc"] + CFG["Nf"]) + 2 * 3 * FLOP_RAY_FWD


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("CODENERF_PRECISION", "bf16"), choices=["bf16", "fp32"])
    ap.add_argument("--cpu-sample-rays", type=int, default=1024)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ddp", action="store_true", help="N > 1: wrap the modules in DistributedDataParallel as "
                    "utils/util.py:139-142 does, instead of codenerf_b200.parallel.GradAllReduce")
    ap.add_argument("--torch-loss", action="store_true", help="2 x torch.nn.functional.mse_loss + add (train.py:103-104 verbatim) "
                    "instead of codenerf_b200.ops.mse_pair")
    ap.add_argument("--no-prefetch", action="store_true", help="e2e leg: draw each step's pixel permutations at the start "
                    "of the step (the reference's order of work) instead of under the previous step's kernels")
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


# ---------------------------------------------------------------------------------------------------------------
def cpu_reference_rays_per_s(n_rays, steps, warmup, seed=0, c=CFG):
    """The reference algorithm on the host cores (oracle port, fp32 torch CPU): fwd + 2 MSE + backward + AdamW on one
    chunk of n_rays rays at C1 shapes.  Returns (rays/s, threads, seconds per step)."""
    from oracle import codenerf_oracle as orc
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    rng = np.random.default_rng(seed)
    cfg = orc.RenderConfig(c["Sc"], c["Nf"], c["near"], c["far"], "lindepth", c["Lx"], c["Ld"])
    pc = {k: v.requires_grad_(True) for k, v in orc.make_codenerf_params(1, c["hidden"], c["code"]).items()}
    pf = {k: v.requires_grad_(True) for k, v in orc.make_codenerf_params(2, c["hidden"], c["code"]).items()}
    tab_s = torch.randn(c["num_objects"], c["code"]).requires_grad_(True)
    tab_t = torch.randn(c["num_objects"], c["code"]).requires_grad_(True)
    opt = torch.optim.AdamW([{"params": list(pc.values())}, {"params": list(pf.values())},
                             {"params": [tab_s, tab_t], "lr": 1e-3}], lr=1e-4)
    imgs, poses, ids = synthetic_batch(rng, c)
    d = orc.camera_directions(c["H"], c["W"], intrinsics(c))
    per = max(1, n_rays // c["B"])
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        sel = orc.draw_select_inds(c["H"] * c["W"], c["B"], per)
        ro, rd = orc.sample_rays(d, torch.from_numpy(poses), sel)
        tgt = torch.cat([torch.from_numpy(imgs).flatten(1, 2)[b, sel[b], :3] for b in range(c["B"])])
        oid = torch.from_numpy(ids)[:, None].expand(-1, per).reshape(-1)
        z_s, z_t = orc.embedding_lookup(tab_s, tab_t, oid)
        rgb_c, rgb_f, _ = orc.predict_radiance_and_render(cfg, ro, rd, pc, pf, z_s, z_t, torch.rand(ro.shape[0], c["Sc"]),
                                                          torch.rand(ro.shape[0], c["Nf"]))
        loss = torch.nn.functional.mse_loss(rgb_c, tgt) + torch.nn.functional.mse_loss(rgb_f, tgt)
        opt.zero_grad()
        loss.backward()
        opt.step()
        _ = loss.item()
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    sec = float(np.mean(times))
    return per * c["B"] / sec, threads, sec


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.cpu_sample_rays
    rps, threads, sec = cpu_reference_rays_per_s(n, args.steps, max(1, min(args.warmup, 1)))
    sample = f"{n} rays of the C2 workload per step (1 chunk, 32+128 samples, fwd+2xMSE+bwd+AdamW), fp32 torch CPU"
    line = {"impl": "reference", "metric": "rays/sec fwd+bwd (coarse+fine)", "value": rps, "unit": "rays/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config("fp32"), "gpu_launches": 0,
            "cpu_baseline": {"value": rps, "unit": "rays/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": rps, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(precision):
    c = CFG
    return {"workload": "srn-cars-code training (SURVEY C2 / BASELINE configs[1]): synthetic 128x128 SRN-car views, "
                        "B=4 images x 4096 rays = 16384 rays per rank per step in 4 chunks of 4096, 32 coarse + 128 fine "
                        "samples, CodeNeRF hidden 256 / codes 256, PE L=10/4, per chunk: embedding lookup + render fwd + "
                        "2 MSE + backward + AdamW step",
            "rays_per_step_per_gpu": c["B"] * c["rays_per_image"], "chunk": c["chunk"], "mlp_precision": precision,
            "pairing": "reference", "loss": "codenerf_b200.ops.mse_pair: both MSE means and their sum in one launch (--torch-loss: "
                                            "2 x torch mse_loss + add)",
            "parallelism": "data-parallel over rays; per chunk step one flat NCCL all-reduce (mean) of all MLP and code-table "
                           "gradients (codenerf_b200.parallel.GradAllReduce; --ddp = DistributedDataParallel wrappers as "
                           "utils/util.py:139-142)",
            "optimizer": "AdamW (lr 1e-4, embeddings 1e-3; train.py:111-114) via codenerf_b200.optim.FusedAdamW, one launch",
            "loss_read": "summed fine loss of the step's 4 chunks, one device->host read per step (e2e leg)",
            "e2e_host_overlap": "prefetching loader: the next step's inputs (pinned H2D copies, np.random.permutation draws from the same "
                                "RNG stream, ray kernel, target gather with the device copy of select_inds) are staged after this "
                                "step's kernels are enqueued and before its loss is read; one staging and one loss read per step",
            "l2": "activations written/read per chunk (>1 GB) exceed the 126 MB L2; no explicit flush"}


# ---------------------------------------------------------------------------------------------------------------
def run_ours(args):
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

    mods = {"embedding": models.ShapeTextureEmbedding(c["num_objects"], c["code"], c["code"]).to(dev),
            "nerf_coarse": models.CodeNeRFModel(c["hidden"], c["num_objects"], c["code"], c["code"], c["Lx"], c["Ld"]).to(dev),
            "nerf_fine": models.CodeNeRFModel(c["hidden"], c["num_objects"], c["code"], c["code"], c["Lx"], c["Ld"]).to(dev)}
    sync = None
    if world > 1 and args.ddp:                                            # utils/util.py:139-142, as the reference wraps them
        from torch.nn.parallel import DistributedDataParallel as DDP
        mods = {k: DDP(m, device_ids=[local], output_device=local) for k, m in mods.items()}
    elif world > 1:
        # the same exchange (mean of every MLP-weight and code-table gradient over ranks, rank 0's initial parameters) as ONE
        # flat NCCL all-reduce per chunk step instead of three DDP reducers: codenerf_b200/parallel.py
        from codenerf_b200.parallel import GradAllReduce
        sync = GradAllReduce([p for m in mods.values() for p in m.parameters()])
    from codenerf_b200.optim import FusedAdamW          # torch.optim.AdamW arithmetic, one launch
    opt = FusedAdamW([{"params": mods["nerf_coarse"].parameters()}, {"params": mods["nerf_fine"].parameters()},
                             {"params": mods["embedding"].parameters(), "lr": 1e-3}], lr=1e-4)
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

    def chunks_step(ro, rd, tgt, oid, read_loss):
        total = None
        for i in range(0, n_rays, c["chunk"]):
            sl = slice(i, i + c["chunk"])
            codes = mods["embedding"](oid[sl])
            rgb_c, rgb_f = nerf.predict_radiance_and_render((ro[sl], rd[sl]), point_sampler, embedders,
                                                            mods["nerf_coarse"], mods["nerf_fine"], codes)
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

    staged = []

    def stage_inputs():
        """One step's host -> device work: pinned images / poses / ids H2D, ray_sampler.sample (host numpy permutation +
        ray kernel), target-pixel gather (train.py:76-80)."""
        imgs = imgs_pin.to(dev, non_blocking=True)
        poses = poses_pin.to(dev, non_blocking=True)
        ids = ids_pin.to(dev, non_blocking=True)
        ro, rd, _sel = ray_sampler.sample(tform_cam2world=poses)         # train.py:76
        sel = ray_sampler.select_inds_device                             # the same indices, already on the device
        flat = imgs.flatten(1, 2)
        tgt = torch.cat([flat[k, sel[k], :] for k in range(c["B"])], dim=0)[..., :3]   # train.py:77-80
        oid = ids[:, None].expand(-1, c["rays_per_image"]).reshape(-1)
        staged.append((ro, rd, tgt, oid))

    def step_e2e():
        # Every step does one stage_inputs() and one loss read.  With prefetching (default) the inputs of step k+1 are staged
        # after step k's kernels have been enqueued and before its loss is read -- a prefetching loader -- so the host
        # permutation and the copies run under the GPU's work; --no-prefetch keeps train.py's order (stage, compute, read).
        if args.no_prefetch or not staged:
            stage_inputs()
        total = chunks_step(*staged.pop(0), True)
        if not args.no_prefetch:
            stage_inputs()
        _ = total.item()                   # the step's result: summed fine loss of its chunks, one device->host read per step

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

    lib = _lib.load()
    step_resident(); torch.cuda.synchronize()
    n0 = lib.cnrf_launch_count(); step_resident(); torch.cuda.synchronize()
    LAUNCHES["per_step"] = lib.cnrf_launch_count() - n0
    clocks = ClockSampler(local)
    clocks.start()
    fine_rows = c["chunk"] * (c["Sc"] + c["Nf"])
    if args.precision == "bf16":
        lib.cnrf_debug_fwd_kernel_ms(fine_rows, None, None)              # time the fine-pass forward-chain launches
    ms_val = timed(step_resident, args.steps, max(3, args.warmup))
    kernel_ms = float(lib.cnrf_debug_fwd_kernel_ms(0, None, None)) if args.precision == "bf16" else 0.0
    clk = clocks.stop()
    ms_e2e = timed(step_e2e, args.steps, 1)
    d2h = 4
    total_rays = n_rays * world * args.steps
    value, e2e = total_rays / (ms_val * 1e-3), total_rays / (ms_e2e * 1e-3)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        roof = kernel_roofline(args, kernel_ms, peaks)
        line = {"metric": "rays/sec fwd+bwd (coarse+fine)", "value": value, "unit": "rays/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_val / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
                "config": dict(workload_config(args.precision), **({"e2e_host_overlap": "off (--no-prefetch)"} if args.no_prefetch else {})),
                "clocks": clk,
                "e2e": {"value": e2e, "unit": "rays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "gpu_launches": roof.pop("launches_per_step"), "roofline": roof,
                "model_flops_per_ray": FLOP_PER_RAY_TRAIN,
                "achieved_tflops_whole_step": value * FLOP_PER_RAY_TRAIN / 1e12 / world}
        line["config"]["gradient_exchange"] = ("none (1 GPU)" if world == 1 else
                                               "DistributedDataParallel x3" if args.ddp else "GradAllReduce (one flat all-reduce)")
        if world == 1:
            # standalone rooflines of the HBM-bound stages (north_star: compositing / sampling vs the HBM peak): one launch
            # per repetition over 524,288 rays (working sets > L2), CUDA events, algorithmic bytes of SURVEY section 8(d)
            import importlib.util
            spec = importlib.util.spec_from_file_location("time_hbm_kernels", os.path.join(ROOT, "scripts", "time_hbm_kernels.py"))
            thk = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(thk)
            line["hbm_kernels"] = [{k: (round(r[k], 4) if isinstance(r[k], float) else r[k]) for k in
                                    ("kernel", "rays", "bytes_per_ray", "ms", "GBps", "frac")}
                                   for r in thk.measure(524288, 10, dev, verbose=False)]
            line["hbm_kernels_peak"] = {"GBps": thk.hbm_peak()[0], "source": thk.hbm_peak()[1]}
            torch.cuda.empty_cache()
        if world == 1 and not args.no_cpu_baseline:
            rps, threads, sec = cpu_reference_rays_per_s(args.cpu_sample_rays, 2, 1)
            line["cpu_baseline"] = {"value": rps, "unit": "rays/s", "cores": threads, "kind": "port",
                                    "sample": f"{args.cpu_sample_rays} rays x 2 timed steps of the same workload "
                                              "(oracle port of the reference path, fp32 torch CPU, all host threads)"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def kernel_roofline(args, kernel_ms, peaks, c=CFG):
    """Dominant kernel = mlp_fwd_tc_kernel on the fine pass (4096 rays x 160 samples through the fused per-sample chain):
    its launches inside the timed region are bracketed by CUDA events on the launching stream by the library itself
    (cnrf_debug_fwd_kernel_ms); achieved = algorithmic FLOPs of one launch / mean duration.  The kernel is timed inside
    a long step, so the denominator is the sustained measured peak."""
    N, S = c["chunk"], c["Sc"] + c["Nf"]
    flops = N * S * FLOP_SAMPLE_FWD
    achieved = flops / (kernel_ms * 1e-3) / 1e12 if kernel_ms > 0 else 0.0
    key = "bf16_tflops_sustained" if "bf16_tflops_sustained" in peaks else "bf16_tflops"
    peak = peaks.get(key, 1400.0)
    traffic = None
    try:   # dram__bytes_read.sum + dram__bytes_write.sum of one fine-pass launch, from the committed ncu --set full capture
        traffic = json.load(open(os.path.join(ROOT, "profiles", "r01_fwd_kernel_ncu.json")))["dram_bytes_per_launch"]
    except Exception:
        pass
    return {"bound": "tensor", "kernel": "mlp_fwd_tc_kernel, fine pass (4096 rays x 160 samples = 655,360 sample rows; "
            "558,592 algorithmic FLOP per sample, DESIGN.md section 4)" if args.precision == "bf16" else "fp32 SIMT chain",
            "achieved": achieved, "peak": peak,
            "peak_source": f"MEASURED_PEAKS.json {key}" if key in peaks else "fallback 1400 TF/s sustained (B200_PROFILING.md)",
            "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic, "ms_per_launch": kernel_ms,
            "algorithmic_flops_per_launch": flops, "launches_per_step": int(LAUNCHES.get("per_step", 0))}


LAUNCHES = {}


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    run_ours(args)


if __name__ == "__main__":
    main()
