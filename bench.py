#!/usr/bin/env python
"""bench.py — SETP1+SETP2 inference frames/s at 480x640 (BASELINE.json metric) on N B200s.

  python bench.py --gpus 1 --steps K --warmup W                 # this repo's CUDA path
  torchrun ... bench.py --gpus N ...                            # one rank per GPU, batch-sharded
  python bench.py --impl reference ...                          # the reference algorithm on host cores

A step = one pass of the hot path (SETP1_NCONV + the RGB-guided SETP2 stack, SETP2_BP_EXPORT
semantics) over one batch of synthetic frames.  Per GPU: 32 frames (16 per stream), i.e.
BASELINE config 4 at N=1; frames are independent, so ranks process disjoint batches with no
data-path collective (weak scaling: global batch = 32 x N).

Printed JSON (rank 0, one line): value = frames/s with inputs resident in HBM; e2e = frames/s
through the public module call with pinned HOST inputs and a HOST result (H2D + D2H inside the
timed region); roofline = the dominant kernel of the step from per-launch CUDA events recorded
by the library on the launching stream during the timed steps; cpu_baseline = the CPU oracle
port (oracle/nconv_oracle.py: the reference's own ATen call sequence) on a bounded sample.

After the headline regions (never inside them) the same line gains:
  parity        the timed batch's frames 0/5/10/15 of both streams against the CPU oracle (max-rel);
  nconv2d_sweep BASELINE config 2: single NConv2d layers, 8 shapes x fwd/bwd at 16x480x640, HBM GB/s vs peak;
  setp1_kitti   BASELINE config 3: SETP1 on 352x1216 KITTI-shape depth, 64 frames split over the N GPUs;
  train_step    BASELINE config 5: SETP1+SETP2 training step, 16 frames per GPU, NCCL all-reduce of the flat
                gradient bucket (the only collective on this path), with its time and bytes.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

METRIC = "setp1+setp2_inference_frames_per_s_480x640"
UNIT = "frames/s"
H, W = 480, 640
FRAMES_PER_STREAM = 16          # x2 streams = 32 frames per GPU (BASELINE config 4)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", os.environ.get("NCONV_B200_SMI_LMS", "100")], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def oracle_model_and_inputs(frames_per_stream: int, seed: int):
    from oracle import nconv_oracle as O
    from nconv_b200.synthetic import setp2_inputs, synthetic_state_dict
    sd = synthetic_state_dict(5)
    return O, sd, setp2_inputs(frames_per_stream, seed=seed, H=H, W=W)


def time_cpu_oracle(frames_per_stream: int, min_seconds: float, max_reps: int, warmup: int, threads: int = 0):
    """The reference's algorithm on the host cores: oracle/nconv_oracle.py issues the same ATen
    calls as /root/reference/models/step1.py + step2.py (eval mode, EXPORT graph)."""
    torch.set_num_threads(threads or os.cpu_count() or 1)
    O, sd, inp = oracle_model_and_inputs(frames_per_stream, seed=7)
    with torch.no_grad():
        for _ in range(warmup):
            O.setp2_forward(sd, *inp, export=True)
        times = []
        t_all = time.perf_counter()
        while len(times) < max_reps and (len(times) < 1 or time.perf_counter() - t_all < min_seconds):
            t0 = time.perf_counter()
            O.setp2_forward(sd, *inp, export=True)
            times.append(time.perf_counter() - t0)
    return times, 2 * frames_per_stream, torch.get_num_threads()


def run_reference(args):
    """--impl reference: rank 0 alone times the CPU implementation; other ranks exit 0."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    fps_frames = 1                                   # 1 frame per stream = 2 frames per step
    torch.set_num_threads(os.cpu_count() or 1)
    O, sd, inp = oracle_model_and_inputs(fps_frames, seed=7)
    with torch.no_grad():
        for _ in range(args.warmup):
            O.setp2_forward(sd, *inp, export=True)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            O.setp2_forward(sd, *inp, export=True)
        dt = time.perf_counter() - t0
    frames = 2 * fps_frames * args.steps
    value = frames / dt
    sample = "%d steps x 2 frames (1 per stream) of the config-4 workload, torch CPU fp32, %d threads" % (
        args.steps, torch.get_num_threads())
    cfg = workload_config(args.gpus)
    cfg["frames_per_step"] = 2 * fps_frames          # the CPU arm's bounded sample of that workload, per step
    cfg["sample"] = "2 of the 32 frames per step (1 per stream); frames are independent, so frames/s is per-frame cost"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(n_gpus: int):
    return {"workload": "BASELINE config 4: SETP1_NCONV + SETP2 (EXPORT graph, eval) inference, synthetic RGB + "
                        "500-point sparse depth at 480x640, 32 frames per GPU (16 per stream)",
            "frames_per_gpu": 2 * FRAMES_PER_STREAM, "global_batch": 2 * FRAMES_PER_STREAM * n_gpus,
            "height": H, "width": W, "parallelism": "batch-sharded x%d, no collective" % n_gpus,
            "l2": "inputs (157 MB/step) and activations (>10 GB/step) exceed the 126 MB L2; no explicit flush"}


def build_model(dev):
    from models.step2 import SETP2_BP_EXPORT
    from nconv_b200.synthetic import synthetic_state_dict     # the same tensors the CPU arm loads
    torch.manual_seed(5)
    m = SETP2_BP_EXPORT()
    missing = m.load_state_dict(synthetic_state_dict(5), strict=False)
    assert not missing.unexpected_keys and all("bnorm" in k for k in missing.missing_keys)
    return m.eval().to(dev)


def summarise_kernels(records, steps):
    by = {}
    for label, ms, fl, by_ in records:
        d = by.setdefault(label, {"ms": 0.0, "n": 0, "flops": 0.0, "bytes": 0.0})
        d["ms"] += ms; d["n"] += 1; d["flops"] += fl; d["bytes"] += by_
    total = sum(d["ms"] for d in by.values()) or 1.0
    rows = []
    for label, d in sorted(by.items(), key=lambda kv: -kv[1]["ms"]):
        avg_ms = d["ms"] / d["n"]
        rows.append({"kernel": label, "launches_per_step": d["n"] / steps, "avg_ms": avg_ms,
                     "share": d["ms"] / total, "flops": d["flops"] / d["n"], "bytes": d["bytes"] / d["n"],
                     "gbs": d["bytes"] / d["n"] / avg_ms / 1e6 if avg_ms > 0 else 0.0,
                     "tflops": d["flops"] / d["n"] / avg_ms / 1e9 if avg_ms > 0 else 0.0})
    return rows


def measured_traffic(kernel_label):
    """DRAM bytes (read + write) of one launch of `kernel_label` from a committed `ncu --set full` capture of this
    very command (profiles/roofline_traffic.json names the capture), or None when no capture covers the kernel."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        with open(path) as f:
            ent = json.load(f).get(kernel_label)
        return float(ent["dram_bytes"]) if ent else None
    except (OSError, ValueError, KeyError, TypeError):
        return None


def roofline_of(row, peaks, sustained=True):
    """Roofline of one kernel from its ALGORITHMIC work: `bound` is whichever of the HBM time (algorithmic bytes /
    measured copy bandwidth) and the tensor time (algorithmic flops / measured bf16 GEMM rate) is longer, `frac` the
    achieved fraction of that bound; `frac_hbm` and `frac_tensor_algorithmic` are always reported side by side.
    The tcgen05 convolutions issue ~3 bf16 MMAs per algorithmic product (bf16x3 split precision: the 1e-4 parity bound
    rules out plain bf16 / TF32); `mma_factor` and `frac_tensor_issued` describe that cost and take no
    part in choosing the bound."""
    tf_peak = peaks["bf16_tflops_sustained" if sustained else "bf16_tflops"]
    tc = row["kernel"].startswith(("tc_conv", "tc_nconv"))
    t_hbm = row["bytes"] / (peaks["hbm_gbs"] * 1e9)
    t_tc = row["flops"] / (tf_peak * 1e12) if tc else 0.0
    frac_hbm = row["gbs"] / peaks["hbm_gbs"]
    out = {"kernel": row["kernel"], "traffic": measured_traffic(row["kernel"]), "peak_source": peaks["source"],
           "avg_ms": row["avg_ms"], "share_of_step": row["share"], "hbm_gbs": row["gbs"], "frac_hbm": frac_hbm,
           "hbm_frac": frac_hbm, "t_hbm_ms": 1e3 * t_hbm}
    if tc:
        mma_factor = 3.0
        ci = int(row["kernel"].split()[1].split("->")[0]) if row["kernel"].startswith("tc_conv3x3 ") else 0
        if ci >= 64 and "->32 " in row["kernel"] and not os.environ.get("NCONV_B200_NO_TCX"):
            # pixels-as-N kernel (tc_conv_x.cuh): per kernel row and 16 input channels three M128 x N256 x K16 MMAs
            # produce 3 taps x 32 channels x 254 pixels (a quarter of the stacked rows is zero, lo*lo rides along)
            mma_factor = (128.0 * 256.0) / (32.0 * 254.0)
        out.update({"tflops": row["tflops"], "frac_tensor_algorithmic": row["tflops"] / tf_peak, "t_tensor_ms": 1e3 * t_tc,
                    "mma_factor": mma_factor, "frac_tensor_issued": mma_factor * row["tflops"] / tf_peak,
                    # the split-precision cost alone (3 bf16 MMAs per algorithmic product, no padding counted): the tensor
                    # time the fp32-level parity bound forces, and the achieved fraction of max(that, HBM time)
                    "t_tensor_bf16x3_ms": 3e3 * t_tc, "frac_of_max_hbm_bf16x3": max(t_hbm, 3.0 * t_tc) / (row["avg_ms"] * 1e-3)})
    if t_hbm >= t_tc:
        out.update({"bound": "hbm", "achieved": row["gbs"], "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": frac_hbm})
        if not tc and row["bytes"] > 0 and row["flops"] / row["bytes"] > 11.0:
            # fp32 CUDA-core kernel above the ~11 flop/B ridge: FMA-bound, HBM fraction reported for reference
            out["note"] = "FMA-bound on CUDA cores (%.0f flop/B): %.1f fp32 TFLOP/s of 72 measured" % (
                row["flops"] / row["bytes"], row["tflops"])
    else:
        out.update({"bound": "tensor", "achieved": row["tflops"], "peak": tf_peak, "unit": "TFLOP/s",
                    "frac": row["tflops"] / tf_peak})
    return out


BENCH_PARITY_FRAMES = (0, 5, 10, 15)


def bench_batch_parity(model, inputs_dev, inputs_host, frames=BENCH_PARITY_FRAMES):
    """Run the config-4 batch through `model` (SETP2_BP_EXPORT) and compare the frames `frames` of BOTH streams with
    the CPU oracle evaluated on exactly those frames (frames are independent).  Checker only, outside every timed
    region; shared with tests/test_gpu_parity.py::test_setp2_bench_batch_parity."""
    from oracle import nconv_oracle as O
    with torch.no_grad():
        o0, o1 = model(*inputs_dev)
    torch.cuda.synchronize()
    idx = torch.tensor(list(frames))
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    sub = [t[idx] for t in inputs_host]
    torch.set_num_threads(os.cpu_count() or 1)
    with torch.no_grad():
        ref = O.setp2_forward(sd, *sub, export=True)
    ours = torch.cat((o0[idx.to(o0.device)], o1[idx.to(o1.device)]), 0).cpu()
    return {"frames_checked": 2 * len(frames), "frame_indices": list(frames), "checker": "oracle/nconv_oracle.py (CPU fp32)",
            "max_rel_floor0p1": O.max_rel_err(ours, ref, floor=0.1), "max_rel_floor1e3": O.max_rel_err(ours, ref, floor=1e-3),
            "max_abs_over_max_ref": float((ours - ref).abs().max() / ref.abs().max())}


# ---- sub-records for BASELINE configs 2, 3, 5 (timed with CUDA events AFTER the headline regions) -------------------
SWEEP_SHAPES = [(ci, co, k) for k in (3, 5) for ci, co in ((1, 2), (2, 2), (4, 4), (16, 16))]


def nconv2d_sweep(dev, peaks, iters=5):
    """Config 2: forward and backward of one NConv2d layer at 16x480x640 through torch.ops.nconv_b200, clean-L2
    protocol of DESIGN 3.1b (256 MB written then read back before each timed launch, device-side spin so the enqueue
    is off the clock), median of `iters`.  Algorithmic bytes as SURVEY 8d."""
    K = torch.ops.nconv_b200
    B = 16
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    sm_max_mhz = float(peaks.get("sm_max_mhz") or 1965.0)
    fma_peak = 148 * 128 * 2 * sm_max_mhz / 1e6          # TFLOP/s: 148 SMs x 128 fp32 lanes x 2 at the maximum SM clock (nominal)

    def timed(fn):
        fn(); fn()
        ts = []
        for _ in range(iters):
            flush.fill_(1); flush.view(torch.float32).sum()
            torch.cuda._sleep(400000)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return sorted(ts)[len(ts) // 2]

    for idx, (ci, co, k) in enumerate(SWEEP_SHAPES):
        pad = k // 2
        g = torch.Generator(device="cpu"); g.manual_seed(1000 + idx)
        x = (torch.rand(B, ci, H, W, generator=g) * 10).to(dev)
        c = (torch.rand(B, ci, H, W, generator=g) * (torch.rand(B, ci, H, W, generator=g) < 0.5)).to(dev)
        w = torch.nn.functional.softplus(torch.rand(co, ci, k, k, generator=g), beta=10).to(dev)
        b = torch.full((co,), 0.01, device=dev)
        gy = torch.randn(B, co, H, W, generator=g).to(dev)
        gc = torch.randn(B, co, H, W, generator=g).to(dev)
        y, co_ = K.nconv2d(x, c, w, b, pad, pad, 1e-7)
        ms_f = timed(lambda: K.nconv2d(x, c, w, b, pad, pad, 1e-7))
        ms_b = timed(lambda: K.nconv2d_backward(gy, gc, x, c, w, b, y, co_, pad, pad, 1e-7, True, True, True))
        bytes_f = 4.0 * B * (2 * ci + 2 * co) * H * W
        bytes_b = 4.0 * B * (4 * co + 4 * ci) * H * W
        flops_f = 4.0 * ci * co * k * k * B * H * W
        # the bounding roofline of a layer is max(t_HBM, t_FMA) -- these kernels are fp32 CUDA-core code, and
        # above Ci*Co = 4 the FMA pipe, not HBM, is the longer of the two (algorithmic flops: fwd 4*Ci*Co*k*k, bwd 8*...)
        t_hbm_f, t_fma_f = bytes_f / peaks["hbm_gbs"] / 1e6, flops_f / fma_peak / 1e9
        t_hbm_b, t_fma_b = bytes_b / peaks["hbm_gbs"] / 1e6, 2 * flops_f / fma_peak / 1e9
        rows.append({"shape": "%d->%d k%d" % (ci, co, k), "fwd_ms": round(ms_f, 4), "fwd_gbs": round(bytes_f / ms_f / 1e6, 1),
                     "fwd_hbm_frac": round(bytes_f / ms_f / 1e6 / peaks["hbm_gbs"], 4), "fwd_tflops": round(flops_f / ms_f / 1e9, 2),
                     "fwd_bound": "fma" if t_fma_f > t_hbm_f else "hbm", "fwd_frac_of_bound": round(max(t_hbm_f, t_fma_f) / ms_f, 4),
                     "bwd_ms": round(ms_b, 4), "bwd_gbs": round(bytes_b / ms_b / 1e6, 1),
                     "bwd_hbm_frac": round(bytes_b / ms_b / 1e6 / peaks["hbm_gbs"], 4),
                     "bwd_tflops": round(2 * flops_f / ms_b / 1e9, 2),
                     "bwd_bound": "fma" if t_fma_b > t_hbm_b else "hbm", "bwd_frac_of_bound": round(max(t_hbm_b, t_fma_b) / ms_b, 4)})
        del x, c, gy, gc, y, co_
    del flush
    return {"workload": "BASELINE config 2: single NConv2d layer, batch 16 at 480x640, fwd and bwd timed separately",
            "protocol": "clean cold L2 (256 MB write + read-back) before every timed launch, median of %d" % iters,
            "hbm_peak_gbs": peaks["hbm_gbs"], "fma_peak_tflops_nominal": round(fma_peak, 1),
            "bound": "per row: the longer of t_HBM (algorithmic bytes / measured copy bandwidth) and t_FMA (algorithmic flops / "
                     "nominal fp32 CUDA-core peak at the maximum clock); frac_of_bound = that time / measured time", "rows": rows}


def setp1_kitti(dev, rank, world, steps, barrier, max_over_ranks):
    """Config 3: SETP1_NCONV inference on KITTI-shape 352x1216 depth (~5 % density), 64 frames split over the ranks
    (strong scaling; 64 per GPU at N=1), no collective.  Layer-wise algorithmic work per frame: SURVEY 8d."""
    from models.step1 import SETP1_NCONV
    from nconv_b200.synthetic import sparse_depth_kitti
    frames = max(64 // world, 1)
    torch.manual_seed(0)
    m = SETP1_NCONV().to(dev)
    m.train()
    with torch.no_grad():
        m(torch.zeros(1, 1, 16, 16, device=dev))       # one training-mode forward: the hook applies softplus
    m.eval()
    S = sparse_depth_kitti(frames, seed=2000 + rank).to(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.no_grad():
        for _ in range(3):
            m(S)
        barrier(); torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            m(S)
        e1.record()
        torch.cuda.synchronize(); barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / steps
    mb, gf = 241.89, 6.569
    return {"workload": "BASELINE config 3: SETP1_NCONV inference, 352x1216, 5 %% density, 64 frames over %d GPU(s)" % world,
            "frames_per_gpu": frames, "value": frames * world / (ms / 1e3), "unit": "frames/s", "ms_per_step": ms, "steps": steps,
            "scaling": "strong", "layerwise_gbs_per_gpu": mb * frames / ms / 1e3, "layerwise_fp32_tflops_per_gpu": gf * frames / ms / 1e3}


def train_step_record(dev, rank, world, steps, barrier, max_over_ranks):
    """Config 5: one SETP1+SETP2 training step = train_step2.py:59-65 (forward of SETP2_BP_TRAIN in train mode,
    multi-resolution masked MSE, backward, AdamW) at 16 frames per GPU, batch-parallel; the gradient sync is ONE NCCL
    all-reduce of the flat fp32 bucket (FusedAdamW.flat_grad), timed separately with CUDA events."""
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import train_bench
    from nconv_b200 import ops
    frames = 16
    step, model, opt, _ = train_bench.make_trainer(dev, frames, H, W, rank)
    for _ in range(2):
        step()
    barrier(); torch.cuda.synchronize()
    n0 = ops.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize(); barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / steps
    launches = (ops.launch_count() - n0) / steps
    ar_ms = None
    if world > 1:
        reps = 10
        for _ in range(2):
            dist.all_reduce(opt.flat_grad)
        barrier(); torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            dist.all_reduce(opt.flat_grad)
        e1.record()
        torch.cuda.synchronize()
        ar_ms = max_over_ranks(e0.elapsed_time(e1)) / reps
    return {"workload": "BASELINE config 5: SETP1+SETP2 training step at 480x640, %d frames per GPU, global batch %d" % (frames, frames * world),
            "value": frames * world / (ms / 1e3), "unit": "frames/s", "ms_per_step": ms, "steps": steps, "scaling": "weak",
            "frames_per_gpu": frames, "global_batch": frames * world, "loss": float(loss), "gpu_launches_per_step": launches,
            "collective": "one NCCL all_reduce(SUM) per step on the flat fp32 gradient bucket" if world > 1 else "none (one GPU)",
            "allreduce_bytes": opt.flat_grad.numel() * 4, "allreduce_ms": ar_ms,
            "allreduce_share_of_step": (ar_ms / ms) if ar_ms else 0.0,
            "peak_mem_gb": torch.cuda.max_memory_allocated(dev) / 2 ** 30}


def run_b200(args):
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from nconv_b200 import _lib, ops
    from nconv_b200.synthetic import setp2_inputs

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    model = build_model(dev)
    host = [t.pin_memory() for t in setp2_inputs(FRAMES_PER_STREAM, seed=7 + rank, H=H, W=W)]
    resident = [t.to(dev) for t in host]
    frames = 2 * FRAMES_PER_STREAM
    h2d = sum(t.numel() * 4 for t in host)
    d2h = 2 * FRAMES_PER_STREAM * H * W * 4

    def step_resident():
        with torch.no_grad():
            return model(*resident)

    from nconv_b200.pipeline import StreamedInference
    streamer = StreamedInference(model, dev)

    def run_e2e(nsteps):
        """Public path: pinned host batches -> module call -> pinned host results, every step; uploads and
        downloads are double-buffered on copy streams (nconv_b200/pipeline.py) so they overlap the kernels."""
        # every step's result lands in pinned host memory; the consumer keeps a strided sample of it
        return streamer.run((host for _ in range(nsteps)), consume=lambda i, outs: [o[..., ::16, ::16].clone() for o in outs])

    for _ in range(max(args.warmup, 3)):
        step_resident()
    torch.cuda.synchronize()

    # ---- timed region: K steps, inputs resident in HBM -------------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    n0 = ops.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier(); torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        step_resident()
    e1.record()
    torch.cuda.synchronize(); barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = ops.launch_count() - n0
    # ---- the same K steps once more with the library's per-launch CUDA events on: kernel list and roofline.  The
    # events (two per launch, ~100 per step) cost ~0.7 ms per step, so they are not recorded inside the region `value`
    # is timed on; `ms_per_step_profiled` says what the step took with them.
    _lib.profile_enable(True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        step_resident()
    e1.record()
    torch.cuda.synchronize(); barrier()
    ms_profiled = e0.elapsed_time(e1)
    records = _lib.profile_records()
    _lib.profile_enable(False)

    # ---- e2e: pinned host inputs -> module call -> host result ------------------------------
    # (diagnostic, untimed: this box's raw pinned-host -> device rate for one batch, so a slow e2e figure can be told
    #  apart from a slow PCIe / NUMA path)
    e0.record()
    for dst, src in zip(resident, host):
        dst.copy_(src, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    h2d_gbs = h2d / e0.elapsed_time(e1) / 1e6
    run_e2e(2)
    barrier(); torch.cuda.synchronize()
    e0.record()
    last = run_e2e(args.steps)                        # returns after the last result is in host memory
    e1.record()
    torch.cuda.synchronize(); barrier()
    assert len(last) == args.steps and all(torch.isfinite(o).all() for outs in last for o in outs)
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop() if rank == 0 else None

    # ---- after the headline regions: parity of the timed batch, then configs 2 / 3 / 5 ---------------
    parity = bench_batch_parity(model, resident, host) if rank == 0 else None
    barrier()
    peaks = measured_peaks()
    del streamer, resident
    model = None
    import gc
    gc.collect(); torch.cuda.empty_cache()
    extra = {}
    if not args.no_extra:
        sub_steps = max(3, min(args.steps, 10))
        extra["setp1_kitti"] = setp1_kitti(dev, rank, world, sub_steps, barrier, max_over_ranks)
        torch.cuda.empty_cache()
        extra["train_step"] = train_step_record(dev, rank, world, max(2, min(args.steps, 5)), barrier, max_over_ranks)
        gc.collect(); torch.cuda.empty_cache()
        if rank == 0:
            extra["nconv2d_sweep"] = nconv2d_sweep(dev, peaks)
        barrier()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    rows = summarise_kernels(records, args.steps)
    roof = roofline_of(rows[0], peaks) if rows else None
    nconv_rows = [r for r in rows if r["kernel"].startswith(("nconv_fwd", "tc_nconv 8", "nconv1_n8"))]
    line = {
        "metric": METRIC, "value": frames * world * args.steps / (ms / 1e3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (fp32 I/O and accumulation; SETP2 dense convs: bf16x3 split operands on tcgen05, fp32 accumulate)",
        "data": "synthetic",
        "config": workload_config(world),
        "e2e": {"value": frames * world * args.steps / (ms_e2e / 1e3), "unit": UNIT,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / args.steps,
                "pinned_h2d_gbs_this_box": h2d_gbs},
        "gpu_launches": int(launches),
        "ms_per_step_profiled": ms_profiled / args.steps,
        "kernel_events": "per-launch CUDA events on the launching stream, from a second pass of the same K steps right after "
                         "the timed region (recording them costs ~0.7 ms per step)",
        "clocks": clocks,
        "roofline": roof,
        "parity": parity, "parity_max_rel": parity["max_rel_floor0p1"] if parity else None,
        "nconv_roofline": [roofline_of(r, peaks) for r in nconv_rows[:4]],
        "kernels": [{k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()} for r in rows[:40]],
    }
    line.update(extra)
    if world == 1 and not args.no_cpu_baseline:
        times, nfr, cores = time_cpu_oracle(1, min_seconds=10.0, max_reps=8, warmup=1)
        best = sorted(times)[len(times) // 2]
        t1, _, _ = time_cpu_oracle(1, min_seconds=0.0, max_reps=2, warmup=0, threads=1)
        line["cpu_baseline"] = {"value": nfr / best, "unit": UNIT, "cores": cores, "kind": "port",
                                "value_1thread": nfr / min(t1),
                                "sample": "%d reps x 2 frames (1 per stream) of the same workload, median; "
                                          "oracle/nconv_oracle.py on torch CPU fp32; value_1thread: best of %d reps "
                                          "with torch.set_num_threads(1)" % (len(times), len(t1))}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the config 2 / 3 / 5 sub-records")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
