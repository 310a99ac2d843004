"""Write a tiny SRN-layout dataset tree (the on-disk format view_synthesis/datasets/dataset.py:60-94 reads) with
procedurally ray-cast objects, so the reference's train.py / eval.py can be run unchanged with no network:

    <root>/srn_cars/cars_{train,val}/<object>/intrinsics.txt        line 1: "f cx cy 0", last line: "H W"
    <root>/srn_cars/cars_{train,val}/<object>/rgb/000000.png ...     128 x 128 RGBA, white (255) background
    <root>/srn_cars/cars_{train,val}/<object>/pose/000000.txt ...    16 floats, camera-to-world, OpenCV axes
                                                                     (the loader right-multiplies diag(1,-1,-1,1), :81)

Every object is an ellipsoid with a differently coloured cap, Lambert-shaded, seen from cameras on a sphere of radius
1.3 looking at the origin (the pose family of eval.py:22-38), so the views of one object are multi-view consistent and
a CodeNeRF trained on them has something to learn.  Test / benchmark infrastructure; needs numpy + PIL only.

    python scripts/make_synthetic_srn.py <root> [--train-objects 6] [--val-objects 2] [--views 10] [--size 128]
"""
import argparse
import os

import numpy as np


def pose_spherical_np(theta, phi, rho):
    """Closed form of eval.py:22-38 (theta = elevation, phi = azimuth), OpenGL axes (-z forward)."""
    c2w = np.eye(4, dtype=np.float64)
    c2w[0, 0], c2w[1, 0] = -np.sin(phi), np.cos(phi)
    c2w[0, 1], c2w[1, 1], c2w[2, 1] = -np.sin(theta) * np.cos(phi), -np.sin(theta) * np.sin(phi), np.cos(theta)
    c2w[0, 2], c2w[1, 2], c2w[2, 2] = np.cos(theta) * np.cos(phi), np.cos(theta) * np.sin(phi), np.sin(theta)
    c2w[:3, 3] = rho * c2w[:3, 2]
    return c2w


def render_view(c2w, size, focal, radii, color, cap_color, cap_axis):
    """Ray-cast one ellipsoid (semi-axes `radii`, centred at the origin) -> (size, size, 4) uint8, background 255."""
    h, w = np.mgrid[0:size, 0:size].astype(np.float64)
    c = size / 2.0
    d_cam = np.stack([(w - c) / focal, -(h - c) / focal, -np.ones_like(w)], axis=-1)         # ray_sampler.py:44-51
    rd = d_cam @ c2w[:3, :3].T
    ro = c2w[:3, 3]
    inv = 1.0 / np.asarray(radii)
    o, d = ro * inv, rd * inv                                                                  # unit-sphere space
    a = (d * d).sum(-1)
    b = 2.0 * (d * o).sum(-1)
    cc = (o * o).sum() - 1.0
    disc = b * b - 4 * a * cc
    hit = disc > 0
    t = np.where(hit, (-b - np.sqrt(np.maximum(disc, 0))) / (2 * a), 0.0)
    hit &= t > 0
    p = ro + rd * t[..., None]
    n = p * inv * inv
    n /= np.maximum(np.linalg.norm(n, axis=-1, keepdims=True), 1e-9)
    light = np.array([0.4, 0.3, 0.86])
    shade = 0.35 + 0.65 * np.clip((n * light).sum(-1), 0, 1)
    is_cap = (p * cap_axis).sum(-1) > 0.35 * np.max(radii)
    base = np.where(is_cap[..., None], np.asarray(cap_color), np.asarray(color))
    rgb = np.clip(base * shade[..., None], 0, 0.97)                                           # never 255: mask = rgb != 255 (:72)
    img = np.full((size, size, 4), 255, dtype=np.uint8)
    img[..., :3] = np.where(hit[..., None], np.round(rgb * 255).astype(np.uint8), 255)
    return img


def write_split(root, split, n_objects, n_views, size, seed):
    from PIL import Image
    rng = np.random.default_rng(seed)
    focal = 131.25 * size / 128.0
    base = os.path.join(root, "srn_cars", f"cars_{split}")
    for obj in range(n_objects):
        odir = os.path.join(base, f"obj{seed:02d}{obj:04d}")
        os.makedirs(os.path.join(odir, "rgb"), exist_ok=True)
        os.makedirs(os.path.join(odir, "pose"), exist_ok=True)
        with open(os.path.join(odir, "intrinsics.txt"), "w") as fh:
            fh.write(f"{focal} {size / 2.0} {size / 2.0} 0.\n0. 0. 0.\n1.\n{size} {size}\n")
        radii = rng.uniform(0.18, 0.42, size=3)
        color, cap_color = rng.uniform(0.15, 0.95, size=3), rng.uniform(0.15, 0.95, size=3)
        cap_axis = np.eye(3)[rng.integers(0, 3)]
        for v in range(n_views):
            theta, phi = rng.uniform(0.15, 1.35), rng.uniform(-np.pi, np.pi)
            c2w_gl = pose_spherical_np(theta, phi, 1.3)
            img = render_view(c2w_gl, size, focal, radii, color, cap_color, cap_axis)
            Image.fromarray(img, mode="RGBA").save(os.path.join(odir, "rgb", f"{v:06d}.png"))
            c2w_cv = c2w_gl @ np.diag([1.0, -1.0, -1.0, 1.0])                                  # undone by dataset.py:81
            np.savetxt(os.path.join(odir, "pose", f"{v:06d}.txt"), c2w_cv.reshape(1, 16))
    return base


def main(root, train_objects=6, val_objects=2, views=10, size=128):
    write_split(root, "train", train_objects, views, size, seed=1)
    write_split(root, "val", val_objects, views, size, seed=2)
    return os.path.join(root, "srn_cars")


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("root")
    ap.add_argument("--train-objects", type=int, default=6)
    ap.add_argument("--val-objects", type=int, default=2)
    ap.add_argument("--views", type=int, default=10)
    ap.add_argument("--size", type=int, default=128)
    a = ap.parse_args()
    print(main(a.root, a.train_objects, a.val_objects, a.views, a.size))
