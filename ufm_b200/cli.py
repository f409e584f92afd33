#!/usr/bin/env python3
"""Headless command line for the engine (replaces ``uniflowmatch/cli.py:85-156``).

    python -m ufm_b200.cli infer SOURCE TARGET [-o DIR] [--model base|refine] [--checkpoint DIR | --config NAME]
    python -m ufm_b200.cli batch PAIRS.txt -o DIR [--batch 16] [...]

``infer`` writes what the reference's ``ufm infer`` writes for the two tensors it computes — ``covisibility_mask.png``
(``(mask * 255).astype(uint8)``, cli.py:141) and a flow colour image ``flow_visualization.png`` — plus the raw
``flow.npy`` (2, H, W) float32 in source-image pixels (the reference's ``flow_vis`` / warp visualisations depend on
packages that are not part of the hot path).  ``batch`` reads a pair list (one ``source target`` per line), groups the
pairs by image size, runs them through ``predict_correspondences_batched`` in pinned host batches and writes one
``<index>_flow.npy`` / ``<index>_covisibility.png`` per pair.  PNG / JPEG decode stays on cv2, as in the reference.

Weights: ``--checkpoint DIR`` (``config.json`` + ``model.safetensors`` as written by the reference's hub mixin, see
``models.from_pretrained``); without it the assumed architecture of ``--config`` is built with seeded random weights
(no hub access offline) and a warning says so.
"""
from __future__ import annotations

import argparse
import os
import sys
from collections import OrderedDict
from typing import Dict, List, Sequence, Tuple


def parse_pair_list(text: str, base: str = "") -> List[Tuple[str, str]]:
    """One ``source target`` per line (whitespace or comma separated); blank lines and ``#`` comments are skipped."""
    pairs = []
    for ln, line in enumerate(text.splitlines(), 1):
        line = line.split("#", 1)[0].strip()
        if not line:
            continue
        parts = line.replace(",", " ").split()
        if len(parts) != 2:
            raise ValueError(f"pair list line {ln}: expected 'source target', got {line!r}")
        pairs.append(tuple(p if os.path.isabs(p) else os.path.join(base, p) for p in parts))
    return pairs


def group_by_shape(shapes: Sequence[Tuple[Tuple[int, int], Tuple[int, int]]], batch: int) -> List[List[int]]:
    """Indices grouped into batches of at most ``batch`` pairs whose (source, target) image sizes agree, in first-seen
    order (a batch is one ``predict_correspondences_batched`` call: all of its images share one H x W per side)."""
    groups: "OrderedDict[Tuple, List[int]]" = OrderedDict()
    for i, s in enumerate(shapes):
        groups.setdefault(tuple(s), []).append(i)
    out = []
    for idx in groups.values():
        out += [idx[k:k + batch] for k in range(0, len(idx), batch)]
    return out


def flow_to_color(flow):
    """(2, H, W) flow -> (H, W, 3) uint8 BGR: hue = direction, value = magnitude relative to the 99th percentile."""
    import cv2
    import numpy as np

    fx, fy = flow[0].astype(np.float32), flow[1].astype(np.float32)
    mag, ang = cv2.cartToPolar(fx, fy, angleInDegrees=True)
    scale = max(float(np.percentile(mag, 99)), 1e-6)
    hsv = np.zeros(fx.shape + (3,), np.uint8)
    hsv[..., 0] = (ang / 2).astype(np.uint8)
    hsv[..., 1] = 255
    hsv[..., 2] = np.clip(mag / scale * 255.0, 0, 255).astype(np.uint8)
    return cv2.cvtColor(hsv, cv2.COLOR_HSV2BGR)


def _load_rgb(path: str):
    import cv2

    img = cv2.imread(path)
    if img is None:
        raise FileNotFoundError(f"could not load image {path!r}")
    return cv2.cvtColor(img, cv2.COLOR_BGR2RGB)


def build_model(args):
    from . import get_config
    from .models import UniFlowMatchClassificationRefinement, UniFlowMatchConfidence

    cls = UniFlowMatchClassificationRefinement if args.model == "refine" else UniFlowMatchConfidence
    if args.checkpoint:
        model = cls.from_pretrained(args.checkpoint)
    else:
        name = args.config or ("ufm_refine_assumed" if args.model == "refine" else "ufm_base_assumed")
        print(f"warning: no --checkpoint given: '{name}' with seeded random-init weights (outputs are not meaningful)",
              file=sys.stderr)
        model = cls(max_pairs_per_call=args.batch, **get_config(name))
    model.max_pairs_per_call = args.batch
    return model.to(f"cuda:{args.device}")


def run_pairs(model, pairs: List[Tuple[str, str]], batch: int):
    """Yields (index, flow (2,H0,W0) float32 numpy, covisibility (H0,W0) float32 numpy or None) for every pair."""
    import numpy as np
    import torch

    imgs = [(_load_rgb(a), _load_rgb(b)) for a, b in pairs]
    shapes = [(s.shape[:2], t.shape[:2]) for s, t in imgs]
    for idx in group_by_shape(shapes, batch):
        src = torch.from_numpy(np.stack([imgs[i][0] for i in idx])).pin_memory()
        tgt = torch.from_numpy(np.stack([imgs[i][1] for i in idx])).pin_memory()
        out = model.predict_correspondences_batched(source_image=src, target_image=tgt)
        flow = out.flow.flow_output.cpu().numpy()
        cov = out.covisibility.mask.cpu().numpy() if out.covisibility is not None else None
        for k, i in enumerate(idx):
            yield i, flow[k], (cov[k] if cov is not None else None)


def _write(out_dir: str, stem: str, flow, cov, single: bool):
    import cv2
    import numpy as np

    os.makedirs(out_dir, exist_ok=True)
    if single:
        np.save(os.path.join(out_dir, "flow.npy"), flow)
        cv2.imwrite(os.path.join(out_dir, "flow_visualization.png"), flow_to_color(flow))
        if cov is not None:
            cv2.imwrite(os.path.join(out_dir, "covisibility_mask.png"), (cov * 255).astype(np.uint8))  # cli.py:141
    else:
        np.save(os.path.join(out_dir, f"{stem}_flow.npy"), flow)
        if cov is not None:
            cv2.imwrite(os.path.join(out_dir, f"{stem}_covisibility.png"), (cov * 255).astype(np.uint8))


def make_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="ufm_b200", description="UFM dense correspondence on the B200 engine (headless)")
    sub = p.add_subparsers(dest="command")

    def common(q):
        q.add_argument("--output", "-o", default=None, help="output directory (default: current directory)")
        q.add_argument("--model", choices=["base", "refine"], default="base")
        q.add_argument("--checkpoint", default=None, help="directory with config.json + model.safetensors")
        q.add_argument("--config", default=None, help="assumed config name when no checkpoint is given")
        q.add_argument("--batch", type=int, default=16, help="pairs per engine call")
        q.add_argument("--device", type=int, default=0)

    q = sub.add_parser("infer", help="one image pair")
    q.add_argument("source")
    q.add_argument("target")
    common(q)
    q = sub.add_parser("batch", help="a list of image pairs")
    q.add_argument("pairs", help="text file: one 'source target' per line")
    common(q)
    return p


def main(argv=None) -> int:
    args = make_parser().parse_args(argv)
    if args.command is None:
        make_parser().print_help()
        return 0
    out_dir = args.output or os.getcwd()
    if args.command == "infer":
        pairs = [(args.source, args.target)]
    else:
        with open(args.pairs) as f:
            pairs = parse_pair_list(f.read(), os.path.dirname(os.path.abspath(args.pairs)))
        if not pairs:
            print("error: empty pair list", file=sys.stderr)
            return 1
    model = build_model(args)
    n = 0
    for i, flow, cov in run_pairs(model, pairs, args.batch):
        _write(out_dir, f"{i:06d}", flow, cov, single=args.command == "infer")
        n += 1
    print(f"{n} pair(s) written to {out_dir}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
