#!/usr/bin/env python
"""bench.py -- image pairs / second of the UFM dense-correspondence forward pass (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl engine|reference] [--workload base16|refine16|hires4|refine16_unet]

A "step" is one pass of the hot path over one batch of synthetic image pairs (random-init weights of the assumed
UFM-Base architecture; iid uint8 images of the named shape):
  * value  : device-timed (CUDA events), inputs already resident in HBM, max over ranks, whole-job pairs/s
  * e2e    : the same metric through the public API `predict_correspondences_batched` with pinned HOST tensors in and
             HOST tensors out (one `ufm_forward_host` C-ABI call per step: H2D + forward + D2H inside the timed region)
  * roofline: the dominant kernel family (the tcgen05 GEMM / implicit-GEMM conv kernel), algorithmic FLOPs / summed
             launch durations measured with CUDA events on the launching stream during the timed steps
  * cpu_baseline: the reference's own CPU fp32 path timed on this box's host cores on a bounded sample (kind
             "reference": the unmodified reference classes from oracle/_ref, staged byte for byte by oracle/build_ref.py;
             kind "port": the oracle restatement, only when no staged reference is present)
  * torch_gpu_baseline (informational): the unmodified reference's eager PyTorch forward on the same B200 under its own
             recipe (bf16 autocast for encoder + info sharing, fp32 / TF32 heads), same batch, CUDA-event timed
`--impl reference` times that CPU path alone, on the engine arm's config; `--impl torch_gpu` the eager GPU forward alone.
Multi-GPU: one process per GPU (torchrun), each rank runs its own batch (weak scaling, no data-path collective).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "image pairs/sec at 560x420 (UFM-Base)"
UNIT = "pairs/s"

WORKLOADS = {
    # name: (config name, inference resolution (W, H), pairs per step, description)
    "base16": ("ufm_base_assumed", (560, 420), 16, "UFM-Base batch 16 pairs 560x420 bf16 on 1xB200 (BASELINE.json configs[1])"),
    "refine16": ("ufm_refine_assumed", (560, 420), 16, "UFM-Refine batch 16 pairs 560x420 (configs[2])"),
    "hires4": ("ufm_base_assumed", (1120, 840), 4, "UFM-Base 1120x840 batch 4 (configs[3])"),
    "refine16_unet": ("ufm_refine_unet", (560, 420), 16, "UFM-Refine + UNet fine-feature branch, batch 16 pairs 560x420 (SURVEY 8f N3)"),
    # inputs at a FOREIGN size: normalise + anti-alias resize in (prep_resize_kernel), nearest / bilinear un-map out (unmap_kernel)
    "foreign16": ("ufm_base_assumed", (560, 420), 16, "UFM-Base batch 16 pairs, 640x480 inputs resized to 560x420 and un-mapped back (SURVEY 8f N1)"),
}
INPUT_HW = {"foreign16": (480, 640)}  # (H0, W0) of the images handed to the API when they differ from the inference resolution


def algorithmic_tflop_per_pair(cfg, refine: bool) -> dict:
    """SURVEY.md section 8(d) formulas: GEMM/layer = 2*M*D^2*(3+1+2r), attention/layer = 4*M*S*D, conv = 2*Cin*Cout*k^2*HW."""
    from ufm_b200.packing import engine_config

    ec = engine_config(cfg)
    hp, wp = ec.img_h // ec.patch, ec.img_w // ec.patch
    n = hp * wp
    S = n + 1
    D, d = ec.enc_dim, ec.info_dim
    enc_gemm = ec.enc_depth * 2 * (2 * S) * D * D * 12 + 2 * (2 * n) * D * 3 * ec.patch ** 2
    enc_attn = ec.enc_depth * 4 * (2 * S) * S * D
    # per token and layer: qkv 3d^2 + proj d^2 + MLP 2 d hidden; cross-attention branches add q d^2 + kv 2d^2 + proj d^2
    per_tok = 4 * d + 2 * ec.info_hidden + (4 * d if ec.info_alternating == 2 else 0)
    info_gemm = ec.info_depth * 2 * (2 * n) * d * per_tok + 2 * (2 * n) * D * d
    info_attn = 0
    for i in range(ec.info_depth):
        if ec.info_alternating == 2:  # self-attention within the view + cross-attention to the other view
            info_attn += 2 * 4 * (2 * n) * n * d
            continue
        span = n if (ec.info_alternating == 1 and i % 2 == 1) else 2 * n
        info_attn += 4 * (2 * n) * span * d
    ld = list(ec.dpt_layer_dims)
    fd, hid = ec.dpt_feature_dim, ec.dpt_hidden
    h3, w3 = (hp - 1) // 2 + 1, (wp - 1) // 2 + 1

    def head(out_ch):
        f = 2 * n * (D * ld[0] + 16 * ld[0] * ld[0]) + 2 * 16 * n * 9 * ld[0] * fd
        f += 2 * n * (d * ld[1] + 4 * ld[1] * ld[1]) + 2 * 4 * n * 9 * ld[1] * fd
        f += 2 * n * d * ld[2] + 2 * n * 9 * ld[2] * fd
        f += 2 * n * d * ld[3] + 2 * h3 * w3 * 9 * ld[3] * ld[3] + 2 * h3 * w3 * 9 * ld[3] * fd
        rcu = lambda px: 2 * 2 * px * 9 * fd * fd
        f += rcu(h3 * w3) + 2 * n * fd * fd                      # refinenet4: RCU2, 1x1 after x2 upsample (cropped)
        f += 2 * rcu(n) + 2 * 4 * n * fd * fd                    # refinenet3
        f += 2 * rcu(4 * n) + 2 * 16 * n * fd * fd               # refinenet2
        f += 2 * rcu(16 * n) + 2 * 64 * n * fd * fd              # refinenet1
        f += 2 * 64 * n * 9 * fd * hid                           # processor conv1 @ 8x
        f += 2 * ec.img_h * ec.img_w * (9 * hid * hid + hid * out_ch)
        return f

    heads = head(ec.flow_out_ch) + (head(ec.unc_out_ch) if ec.unc_out_ch else 0)
    cls = 0
    if refine:
        cls = 2 * (2 * n) * (2 * ec.cls_in * ec.cls_hidden + ec.cls_in * ec.cls_out * ec.cls_patch ** 2)
    unet = 0
    if ec.unet:  # unet_encoder.py:26-72 on both images + final / combine 1x1 convs
        F = (64, 128, 256, 512)
        lh, lw, cin = ec.img_h, ec.img_w, 3
        dims = []
        for f in F:
            unet += 2 * lh * lw * 9 * (cin * f + f * f)
            dims.append((lh, lw))
            lh, lw, cin = lh // 2, lw // 2, f
        unet += 2 * lh * lw * 9 * (cin * 2 * cin + 4 * cin * cin)
        for f, (sh, sw) in zip(reversed(F), reversed(dims)):
            unet += 2 * lh * lw * 4 * (2 * f) * f + 2 * sh * sw * 9 * (2 * f * f + f * f)
            lh, lw = sh, sw
        unet += 2 * ec.img_h * ec.img_w * (64 * 16 + (32 * 32 + 32 * 16 if ec.unet == 1 else 16 * 16))
        unet *= 2
    t = {"enc_gemm": enc_gemm, "enc_attn": enc_attn, "info_gemm": info_gemm, "info_attn": info_attn, "heads": heads,
         "cls_head": cls + unet}
    t = {k: v / 1e12 for k, v in t.items()}
    t["total"] = sum(t.values())
    t["gemm_conv"] = t["enc_gemm"] + t["info_gemm"] + t["heads"] + t["cls_head"]
    return t


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md clocks line)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self) -> dict:
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2])); pw.append(float(c[3]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


def _peaks() -> dict:
    """MEASURED_PEAKS.json (driver-written) else the fallback B200_PROFILING.md states."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"bf16_burst": d.get("bf16_tflops", 1590.0), "bf16_sustained": d.get("bf16_tflops_sustained", 1400.0),
                "hbm": d.get("hbm_gbs", 6650.0), "source": "measured"}
    return {"bf16_burst": 1590.0, "bf16_sustained": 1400.0, "hbm": 6650.0, "source": "fallback"}


def _ncu_traffic(workload: str):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` summary (profiles/*_traffic.json,
    written when a capture is summarised); None when no capture of this workload is committed."""
    import glob

    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")), reverse=True):
        try:
            with open(path) as f:
                d = json.load(f)
            if d.get("workload") == workload:
                return d["dram_bytes_per_launch"]
        except Exception:
            continue
    return None


def _make_cfg(workload: str):
    from ufm_b200 import configs

    name, res, pairs, desc = WORKLOADS[workload]
    cfg = configs.CONFIGS[name]() if name == "ufm_refine_unet" else getattr(configs, name)(res)
    return cfg, pairs, desc, name


def _config_dict(desc, pairs, W, H, world):
    """`config` of the JSON line -- the same dict for the engine and the reference arm (the reference arm times a
    bounded sample of this workload and says so in `cpu_baseline.sample`)."""
    return {"workload": desc, "pairs_per_step_per_gpu": pairs, "resolution": [W, H], "weights": "random-init seed 0",
            "l2": "no flush needed: per-step working set (0.9 GB weights + >10 GB activations) exceeds the 126 MB L2",
            "parallelism": f"dp{world} (independent pairs, no data-path collective)"}


def _reference_callable(cfg, device="cpu"):
    """(callable(src, tgt) -> result, kind).  kind "reference": the UNMODIFIED reference classes (oracle/_ref staged by
    oracle/build_ref.py, or /root/reference in the build container) with the seed-0 state dict; kind "port": the
    oracle restatement (CPU only), used when no reference tree is present."""
    import torch

    from ufm_b200.weights import make_state_dict

    sd = make_state_dict(cfg, seed=0)
    from oracle import ref_harness as R

    if R.reference_available():
        model = R.build_reference(cfg, sd).to(device)

        def call(src, tgt):
            with torch.no_grad():
                return model.predict_correspondences_batched(src, tgt)

        return call, "reference"
    if device != "cpu":
        raise SystemExit("bench.py: no staged reference (oracle/_ref) for the torch GPU baseline")
    from oracle import ufm_oracle as O

    def call(src, tgt):
        with torch.no_grad():
            return O.predict_correspondences_batched(src, tgt, sd, cfg)

    return call, "port"


def cpu_reference_pairs_per_s(cfg, n_calls: int = 1, warmup: int = 1, pairs_per_call: int = 1):
    """The reference's CPU fp32 path on this box's host cores, `pairs_per_call` pairs per call (a bounded sample of the
    workload).  Returns (pairs/s, threads, seconds per call list, kind)."""
    import torch

    from oracle.golden_util import make_inputs

    threads = len(os.sched_getaffinity(0))
    torch.set_num_threads(threads)
    call, kind = _reference_callable(cfg, "cpu")
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(pairs_per_call, H, W, seed=1)
    ts = []
    for i in range(warmup + n_calls):
        t0 = time.perf_counter()
        call(src, tgt)
        dt = time.perf_counter() - t0
        if i >= warmup:
            ts.append(dt)
    return pairs_per_call / statistics.median(ts), threads, ts, kind


def run_reference(args):
    """--impl reference: the reference's own CPU fp32 implementation of the path (unmodified classes, all host threads)
    for the engine arm's metric and config; each step is a bounded sample of the workload (one pair per call)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return 0
    cfg, pairs, desc, name = _make_cfg(args.workload)
    if args.pairs:
        pairs = args.pairs
    t0 = time.perf_counter()
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    v, threads, ts, kind = cpu_reference_pairs_per_s(cfg, n_calls=steps, warmup=warmup)
    Wr, Hr = cfg["inference_resolution"]
    what = "unmodified reference classes (oracle/_ref)" if kind == "reference" else "oracle port of the reference path"
    sample = (f"{len(ts)} timed step(s) after {warmup} warm-up, each ONE {Wr}x{Hr} pair (batch 1) of the workload's "
              f"{pairs}-pair batch, CPU fp32, {threads} threads, {what}")
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": len(ts),
        "warmup": warmup, "ms_per_step": 1e3 * statistics.median(ts), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": _config_dict(desc, pairs, Wr, Hr, world),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t0,
    }
    print(json.dumps(line))
    return 0


def run_torch_gpu(args):
    """--impl torch_gpu: the UNMODIFIED reference's eager PyTorch forward on one B200 under its own recipe
    (ref base.py:272-274 bf16 autocast; ufm.py:635 / :921 fp32 heads; TF32 matmul as transformer_blocks.py:20), same
    batch as the engine arm, device-resident uint8 inputs, CUDA-event timed.  Informational: the same-box PyTorch
    number BASELINE.md section 2 names.  No engine code runs in this process."""
    import torch

    from oracle.golden_util import make_inputs

    cfg, pairs, desc, name = _make_cfg(args.workload)
    if args.pairs:
        pairs = args.pairs
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    call, kind = _reference_callable(cfg, dev)
    W, H = cfg["inference_resolution"]
    src, tgt = make_inputs(pairs, H, W, seed=1)
    src, tgt = src.to(dev), tgt.to(dev)
    for _ in range(max(2, args.warmup)):
        call(src, tgt)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        out = call(src, tgt)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    line = {"impl": "torch_gpu", "metric": METRIC, "value": pairs / (ms / 1e3), "unit": UNIT, "ms_per_step": ms,
            "steps": args.steps, "warmup": max(2, args.warmup), "dtype": "bf16 autocast + fp32/TF32 heads",
            "config": _config_dict(desc, pairs, W, H, 1), "kind": kind,
            "what": "unmodified reference, eager PyTorch (cuBLAS / cuDNN / SDPA), predict_correspondences_batched on device uint8",
            "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9}
    print(json.dumps(line))
    return 0


def _torch_gpu_baseline_subprocess(args, local: int):
    """Run `--impl torch_gpu` in a fresh process (the engine's process never loads the reference or its libraries)."""
    from oracle import ref_harness as R

    if not R.reference_available():
        return {"unavailable": "no staged reference (oracle/_ref) on this box"}
    env = dict(os.environ, LOCAL_RANK=str(local))
    for k in ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "torch_gpu", "--workload", args.workload, "--steps",
           str(max(3, min(args.steps, 5))), "--warmup", "2"] + (["--pairs", str(args.pairs)] if args.pairs else [])
    try:
        p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
        for ln in reversed(p.stdout.strip().splitlines()):
            if ln.startswith("{"):
                d = json.loads(ln)
                return {k: d[k] for k in ("value", "unit", "ms_per_step", "steps", "dtype", "what", "kind", "peak_mem_gb")}
        return {"unavailable": f"exit {p.returncode}: {p.stderr.strip().splitlines()[-1] if p.stderr.strip() else 'no output'}"}
    except Exception as e:  # noqa: BLE001
        return {"unavailable": repr(e)}


def run_engine(args):
    import torch
    import torch.distributed as dist

    import __graft_entry__ as g

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (ufm_b200 has no CPU path; use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # keep stdout for the one JSON line: libraries (NCCL's version banner) write to fd 1 during init
    sys.stdout.flush()
    saved_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    g.build()
    from oracle.golden_util import make_inputs
    from ufm_b200 import UniFlowMatchClassificationRefinement, UniFlowMatchConfidence, _lib
    from ufm_b200.weights import make_state_dict

    cfg, pairs, desc, name = _make_cfg(args.workload)
    if args.pairs:
        pairs = args.pairs
    refine = "classification_head_kwargs" in cfg
    W, H = cfg["inference_resolution"]
    cls = UniFlowMatchClassificationRefinement if refine else UniFlowMatchConfidence
    model = cls(max_pairs_per_call=pairs, **cfg)
    model.load_state_dict(make_state_dict(cfg, seed=0))
    model = model.to(dev)
    H0, W0 = INPUT_HW.get(args.workload, (H, W))
    foreign = (H0, W0) != (H, W)
    src, tgt = make_inputs(pairs, H0, W0, seed=1 + rank)
    dsrc, dtgt = src.to(dev), tgt.to(dev)
    eng = model._engine(dev)

    class _Api:  # foreign input size: the timed call is the public API on device tensors (resize in, forward, un-map out)
        def forward(self, a, b, layout):
            o = model.predict_correspondences_batched(a, b)
            return o.flow.flow_output, o.covisibility.mask

        def launch_count(self):
            return eng.launch_count() + 3  # + 2 prep_resize + 1 unmap launches around the engine's pass

    run = _Api() if foreign else eng

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def maxreduce(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    # ---- value: device-timed, inputs resident in HBM -----------------------------------------------------------------
    for _ in range(args.warmup):
        run.forward(dsrc, dtgt, _lib.UFM_NHWC)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    e0.record()
    for i in range(args.steps):
        flow, cov = run.forward(dsrc, dtgt, _lib.UFM_NHWC)
        marks[i].record()
    e1.record()
    barrier()
    clocks = sampler.stop()
    step_ms = [round((e0 if i == 0 else marks[i - 1]).elapsed_time(marks[i]), 3) for i in range(args.steps)]
    launches = run.launch_count() * args.steps
    ms = maxreduce(e0.elapsed_time(e1))
    ms_per_step = ms / args.steps
    value = world * pairs * args.steps / (ms / 1e3)

    # ---- e2e: public API, pinned host tensors in, host tensors out ---------------------------------------------------
    hsrc, htgt = src.pin_memory(), tgt.pin_memory()
    # warm-up calls keep their outputs alive so that the pinned-host allocator's cache holds the two result blocks the
    # timed loop alternates between (a first-use cudaHostAlloc of 45 MB costs 10-20 ms)
    keep = [model.predict_correspondences_batched(hsrc, htgt) for _ in range(max(2, min(args.warmup, 3)))]
    del keep
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        out = model.predict_correspondences_batched(hsrc, htgt)
    e1.record()
    barrier()
    e2e_ms = maxreduce(max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3))
    assert out.flow.flow_output.device.type == "cpu"
    h2d = 2 * pairs * H0 * W0 * 3
    d2h = pairs * H0 * W0 * 4 * (2 + (1 if out.covisibility is not None else 0))
    e2e = {"value": world * pairs * args.steps / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "api": "predict_correspondences_batched(pinned host uint8 BHWC) -> host tensors"}

    # ---- roofline: per-kernel-family durations (CUDA events around every launch, on the launching stream) ------------
    eng.set_profile(True)
    psteps = max(1, min(args.steps, 3))
    for _ in range(psteps):
        run.forward(dsrc, dtgt, _lib.UFM_NHWC)
    prof = eng.profile_read()
    eng.set_profile(False)
    prepost_fam = {}
    if foreign:
        # the two kernels around the engine's pass run outside its launch wrapper: timed here, back to back, on the
        # current stream; algorithmic bytes = input read once + output written once
        from ufm_b200 import prepost

        def timed(fn, reps=10):
            for _ in range(2):
                fn()
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a_.record()
            for _ in range(reps):
                fn()
            b_.record()
            torch.cuda.synchronize()
            return a_.elapsed_time(b_) / reps

        t_prep = timed(lambda: prepost.prep_resize(dsrc, True, H, W, "dinov2"))
        rep0 = prepost.scaled_region(H, W, H0, W0)
        full = [0, H0, 0, W0]
        f_in = torch.randn(pairs, 2, H, W, device=dev)
        c_in = torch.rand(pairs, H, W, device=dev)
        t_unmap = timed(lambda: prepost.unmap(f_in, c_in, rep0, full, full, (H0, W0)))
        b_prep = pairs * (H0 * W0 * 3 + H * W * 3 * 4)
        b_unmap = pairs * 3 * 4 * (H * W + H0 * W0)
        peaks_ = _peaks()
        for nm, t_, by_, n_ in (("prep_resize", t_prep, b_prep, 2), ("unmap", t_unmap, b_unmap, 1)):
            prepost_fam[nm] = {"ms_per_step": t_ * n_, "launches_per_step": n_, "bound": "hbm", "achieved": by_ / (t_ * 1e-3) / 1e9,
                               "peak": peaks_["hbm"], "unit": "GB/s", "frac": by_ / (t_ * 1e-3) / 1e9 / peaks_["hbm"],
                               "note": "timed alone (kernel-level C-ABI call), algorithmic bytes = input once + output once"}
    peaks = _peaks()
    fl = algorithmic_tflop_per_pair(cfg, refine)
    fam = {}
    tot_ms = sum(v["ms"] for v in prof.values())
    for k, v in prof.items():
        if not v["launches"]:
            continue
        d = {"ms_per_step": v["ms"] / psteps, "launches_per_step": v["launches"] // psteps, "share": v["ms"] / tot_ms}
        # SIMT families carry flops only for the UNet stem (K = 27, no tensor-core shape) and the per-pixel combine: they
        # are judged by bytes like the other copies, not against the tensor peak
        if v["flops"] > 0 and k not in ("move", "layernorm", "resize", "refine"):
            d.update(bound="tensor", achieved=v["flops"] / (v["ms"] * 1e-3) / 1e12, peak=peaks["bf16_sustained"], unit="TFLOP/s")
        else:
            d.update(bound="hbm", achieved=v["bytes"] / (v["ms"] * 1e-3) / 1e9, peak=peaks["hbm"], unit="GB/s")
        d["frac"] = d["achieved"] / d["peak"]
        fam[k] = d
    fam.update(prepost_fam)
    gm = {k: prof[k] for k in ("gemm", "conv3x3")}
    g_ms = sum(v["ms"] for v in gm.values())
    g_fl = sum(v["flops"] for v in gm.values())
    g_n = sum(v["launches"] for v in gm.values())
    roofline = {"kernel": "gemm_kernel<BN> (tcgen05 GEMM + implicit-GEMM 3x3 conv)", "bound": "tensor",
                "achieved": g_fl / (g_ms * 1e-3) / 1e12, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                "peak_source": f"{peaks['source']} bf16 sustained", "traffic": _ncu_traffic(args.workload),
                "avg_launch_ms": g_ms / g_n, "flops_per_launch": g_fl / g_n, "share_of_step": g_ms / tot_ms}
    roofline["frac"] = roofline["achieved"] / roofline["peak"]
    whole = {"tflop_per_pair": fl["total"], "achieved_tflops_per_gpu": value / world * fl["total"],
             "frac_of_bf16_sustained": value / world * fl["total"] / peaks["bf16_sustained"]}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": _config_dict(desc, pairs, W, H, world),
        "step_ms": step_ms, "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "kernel_families": fam,
        "whole_pass": whole,
    }
    if rank == 0 and world == 1 and not args.no_torch_gpu:
        # a fresh process: the engine's process never imports the reference or loads cuBLAS / cuDNN
        tg = _torch_gpu_baseline_subprocess(args, local)
        line["torch_gpu_baseline"] = tg
        if "value" in tg:
            line["vs_torch_gpu"] = value / tg["value"]
    if rank == 0 and world == 1 and not args.no_cpu:
        v, threads, ts, kind = cpu_reference_pairs_per_s(cfg, n_calls=args.cpu_calls, warmup=1)
        what = "unmodified reference classes (oracle/_ref)" if kind == "reference" else "oracle port of the reference path"
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": kind,
                                "sample": f"{len(ts)} timed call(s) after 1 warm-up, each ONE {W}x{H} pair (batch 1) of the "
                                          f"workload's {pairs}-pair batch, CPU fp32, {what}"}
    sys.stdout.flush()
    os.dup2(saved_fd, 1)
    if rank == 0:
        print(json.dumps(line), flush=True)
    os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference", "torch_gpu"])
    ap.add_argument("--workload", default="base16", choices=list(WORKLOADS))
    ap.add_argument("--pairs", type=int, default=0, help="override pairs per step per GPU")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-torch-gpu", action="store_true", help="skip the informational eager-PyTorch GPU baseline")
    ap.add_argument("--cpu-calls", type=int, default=3)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "engine" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    if args.impl == "torch_gpu":
        return run_torch_gpu(args)
    return run_engine(args)


if __name__ == "__main__":
    sys.exit(main())
