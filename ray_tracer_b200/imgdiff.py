"""Compare two rendered images: python -m ray_tracer_b200.imgdiff REF.png TEST.png [--map OUT.png] [--ppd N]

RMSE over the sRGB-encoded RGB values and mean LDR-FLIP (host/rt_imgdiff.cpp), the measures of the
tolerance tiers in SURVEY.md §8(c).  Inputs are PNG files as Context.save_png writes them, or .npy
RGBA32F accumulation images as RayTracer.read_accum returns them (tone-mapped here: clamp + sRGB OETF,
what render.wgsl:34-39 and the sRGB swapchain show).  Prints one JSON line.
"""
import argparse
import json

import numpy as np

from . import host


def _load(path):
    if path.endswith(".npy"):
        return host.tonemap(np.load(path))
    return host.load_png(path)[..., :3].astype(np.float32) / np.float32(255)


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m ray_tracer_b200.imgdiff")
    ap.add_argument("ref")
    ap.add_argument("test")
    ap.add_argument("--map", help="write the per-pixel FLIP error as a grey PNG")
    ap.add_argument("--ppd", type=float, default=host.FLIP_PPD, help="pixels per degree (default: the paper's 0.7 m / 3840 px / 0.7 m)")
    a = ap.parse_args(argv)
    r, t = _load(a.ref), _load(a.test)
    if r.shape != t.shape:
        raise SystemExit("imgdiff: image sizes differ: %s vs %s" % (r.shape, t.shape))
    mean, emap = host.image_flip(r, t, a.ppd, want_map=True)
    out = {"width": int(r.shape[1]), "height": int(r.shape[0]), "rmse": host.image_rmse(r, t), "flip_mean": mean,
           "flip_max": float(emap.max()), "differing_pixels": int(np.any(r != t, axis=-1).sum()), "ppd": a.ppd}
    if a.map:
        g = np.clip(emap * 255.0 + 0.5, 0, 255).astype(np.uint8)
        host.write_png(a.map, np.stack([g, g, g, np.full_like(g, 255)], -1))
    print(json.dumps(out))
    return out


if __name__ == "__main__":
    main()
