"""Stage the UNMODIFIED reference for the GPU box: ``python -m oracle.build_ref``  (TEST INFRASTRUCTURE).

``/root/reference`` exists only in the build container.  This recipe copies, byte for byte, exactly the reference
source files that the path of SURVEY.md section 8(a) imports -- found by constructing the reference models of every
info-sharing / head / UNet variant under :mod:`oracle.ref_harness` and reading ``sys.modules`` -- into
``oracle/_ref/`` (git-ignored, NOT gpurun-ignored: it travels to the GPU box like the built ``.so``).  Nothing is
edited; ``oracle/_ref/MANIFEST.txt`` records every file with its sha256 so a reader can check that.  On the GPU box
:mod:`oracle.ref_harness` falls back to this copy, which gives
  * ``bench.py --impl reference``: the reference's own CPU fp32 path timed on the box's host cores (kind "reference"),
  * ``bench.py``'s ``torch_gpu_baseline``: the reference's eager forward on the B200 under its own autocast recipe,
  * ``tests/test_gpu_vs_reference.py``: engine vs reference-on-GPU parity at the full architecture sizes.
Reference sources never enter the git history.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference"
DST = os.path.join(ROOT, "oracle", "_ref")
# one config per code path of the reference that the engine covers (encoder, three info-sharing transformers,
# DPT heads with / without BatchNorm, Refine, UNet branch, RoPE, covariance / confidence adaptors)
PROBE_CONFIGS = ["tiny_base", "tiny_refine", "tiny_cross", "tiny_unet", "tiny_rope", "tiny_cov", "tiny_bn"]


def _closure():
    sys.path.insert(0, ROOT)
    os.environ["UFM_REFERENCE_ROOT"] = SRC
    from oracle.golden_util import case_config
    from oracle.ref_harness import build_reference
    from ufm_b200.configs import tiny

    for name in PROBE_CONFIGS:
        build_reference(case_config(name))
    build_reference(tiny(True, info_sharing_str="alternating_attention", qk_norm=True))
    files = set()
    for k, m in list(sys.modules.items()):
        f = getattr(m, "__file__", None)
        if f and os.path.abspath(f).startswith(SRC + os.sep):
            files.add(os.path.abspath(f))
    # package markers on the way down (namespace packages have none)
    for f in list(files):
        d = os.path.dirname(f)
        while d.startswith(SRC + os.sep):
            init = os.path.join(d, "__init__.py")
            if os.path.exists(init):
                files.add(init)
            d = os.path.dirname(d)
    return sorted(files)


def stage(force: bool = False) -> str | None:
    """Copy the import closure into oracle/_ref; returns the destination (None when /root/reference is absent)."""
    if not os.path.isdir(os.path.join(SRC, "uniflowmatch")):
        return None
    manifest = os.path.join(DST, "MANIFEST.txt")
    if os.path.exists(manifest) and not force:
        return DST
    files = _closure()
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    lines = []
    for f in files:
        rel = os.path.relpath(f, SRC)
        out = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        shutil.copyfile(f, out)
        with open(f, "rb") as fh:
            lines.append(f"{hashlib.sha256(fh.read()).hexdigest()}  {rel}")
    with open(manifest, "w") as fh:
        fh.write("# byte-for-byte copies from /root/reference (labrat97/UFM), made by oracle/build_ref.py\n")
        fh.write("\n".join(lines) + "\n")
    return DST


if __name__ == "__main__":
    d = stage(force="--force" in sys.argv)
    print(f"staged reference at {d}" if d else "no /root/reference here: nothing staged")
