#!/usr/bin/env python
"""bench.py — throughput of the WeCLIP hot path on B200 (contract: see the task statement / DESIGN.md §Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One JSON line on stdout (rank 0).  A "step" = one pass of the hot path over one batch of synthetic input.
N>1 is launched by torchrun (one rank per GPU); images are sharded across ranks with no data-path
collective (weak scaling); the timed region is bracketed by barrier + synchronize and the max over ranks
is taken.  `--impl reference` times the CPU oracle port (the reference is Python and cannot travel to the
GPU box; oracle/weclip_oracle.py is its pinned restatement) on the host cores, one dataset image per step, honouring
--steps / --warmup as given.

Keys beyond the base contract (msc+flip workloads):
  value / e2e           the reference's call pattern: names = name+name with every [x, flip(x)] call
                        (test_msc_flip_voc.py:52,69,84) -> config.cam_passes_per_image = 6
  value_live_outputs    the same step with the passes whose labels the driver never reads eliminated (1 pass; identical
                        histograms, tests/test_evaluate_gpu.py::test_cam_passes_live_equals_reference_call_pattern)
  gpu_library_baseline  the same loop as ATen/cuBLAS library calls on this GPU in the reference's mixed precision (N=1 only)
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

import torch  # noqa: E402

METRIC = "msc+flip CAM/seg inference images/sec"


def _peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.idx)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except ValueError:
                continue
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------
# Workloads
# ----------------------------------------------------------------------------------------------------
class ParPseudoLabel:
    """Interim round-1 workload: the PAR + argmax/LUT tail of BASELINE.json configs[1] (batch 32 crops of
    320x320, K+1 = 3 CAM channels per image, dilations 1/2/4/8/12/24, 20 iterations)."""
    name = "cfg2-tail: PAR(20 it, 6 dil, fp16 affinity)+argmax on 32 x 320x320 crops, K+1=3"
    aff_dtype, s_a = torch.float16, 2
    units = 32
    l2_note = "inputs larger than L2 (affinity 315 MB fp16 / 629 MB fp32 per step)"

    def __init__(self, rank: int, device):
        from weclip_b200 import synth
        from weclip_b200.par import PAR
        self.B, self.H, self.W, self.Kc = 32, 320, 320, 3
        self.dev = device
        g = torch.Generator().manual_seed(100 + rank)
        self.imgs_h = synth.make_images(200 + rank, self.B, self.H, self.W).pin_memory()
        cams = torch.rand(self.B, self.Kc - 1, self.H, self.W, generator=g)
        bg = 1 - cams.max(1, keepdim=True)[0]
        self.masks_h = torch.cat([bg, cams], 1).reshape(self.B * self.Kc, self.H, self.W).contiguous().pin_memory()
        self.ch_off = [b * self.Kc for b in range(self.B + 1)]
        self.lut = torch.tensor([0, 3, 7] * self.B, dtype=torch.int64, device=device)
        self.par = PAR([1, 2, 4, 8, 12, 24], 20, affinity_dtype=self.aff_dtype).to(device)
        self.imgs_d = self.imgs_h.to(device)
        self.masks_d = self.masks_h.to(device)
        self.labels_h = torch.empty(self.B, self.H, self.W, dtype=torch.int64).pin_memory()
        self.launches = 1 + 20 + self.B           # affinity + 20 propagation + per-image argmax
        self.h2d = self.imgs_h.numel() * 4 + self.masks_h.numel() * 4
        self.d2h = self.labels_h.numel() * 8
        self.kernel_events = []

    def step_device(self, record=False):
        from weclip_b200.par import argmax_lut
        if record:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        out = self.par.refine_ragged(self.imgs_d, self.masks_d, self.ch_off)
        if record:
            e1.record()
            self.kernel_events.append((e0, e1))
        return argmax_lut(out, self.lut, self.ch_off)

    def step_e2e(self):
        from weclip_b200.par import argmax_lut
        imgs = self.imgs_h.to(self.dev, non_blocking=True)
        masks = self.masks_h.to(self.dev, non_blocking=True)
        out = self.par.refine_ragged(imgs, masks, self.ch_off)
        lab = argmax_lut(out, self.lut, self.ch_off)
        self.labels_h.copy_(lab, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.labels_h

    def roofline(self, peaks):
        """Dominant kernel: par_iter_tile_kernel.  Algorithmic bytes per launch = H*W*B*(48*s_a + 2*4*Kc) (SURVEY §8d,
        s_a = bytes per stored affinity); the bracketed region holds 1 affinity launch (12+48*s_a B/px) + 20 iteration
        launches."""
        px = self.B * self.H * self.W
        alg = px * (12 + 48 * self.s_a) + 20 * px * (48 * self.s_a + 8 * self.Kc)
        ms = statistics.mean(a.elapsed_time(b) for a, b in self.kernel_events)
        ach = alg / (ms * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": "par_aff_tile_kernel + 20 x par_iter_tile_kernel", "achieved": round(ach, 1),
                "peak": peaks["hbm_gbs"], "peak_source": peaks["source"] + " (burst copy)", "unit": "GB/s",
                "frac": round(ach / peaks["hbm_gbs"], 4), "traffic": None,
                "algorithmic_bytes_per_step": alg, "kernel_ms_per_step": round(ms, 4)}

    @classmethod
    def cpu(cls):
        """CPU-only instance for the reference arm (no CUDA objects)."""
        from weclip_b200 import synth
        wl = cls.__new__(cls)
        wl.B, wl.H, wl.W, wl.Kc = 4, 320, 320, 3
        g = torch.Generator().manual_seed(100)
        wl.imgs_h = synth.make_images(200, wl.B, wl.H, wl.W)
        cams = torch.rand(wl.B, wl.Kc - 1, wl.H, wl.W, generator=g)
        wl.masks_h = torch.cat([1 - cams.max(1, keepdim=True)[0], cams], 1).reshape(wl.B * wl.Kc, wl.H, wl.W)
        return wl

    # CPU baseline (oracle port of PAR.forward), bounded sample
    def cpu_sample(self, n_img=2):
        from oracle import weclip_oracle as O
        imgs = self.imgs_h[:n_img].clone()
        masks = self.masks_h.reshape(self.B, self.Kc, self.H, self.W)[:n_img].clone()
        t0 = time.perf_counter()
        for i in range(n_img):
            out = O.par_forward(imgs[i:i + 1], masks[i:i + 1])
            out.argmax(1)
        dt = time.perf_counter() - t0
        return n_img / dt, f"{n_img} of the step's {self.B} images (PAR+argmax), torch CPU fp32"


class ParPseudoLabelF32(ParPseudoLabel):
    name = "cfg2-tail: PAR(20 it, 6 dil, fp32 affinity)+argmax on 32 x 320x320 crops, K+1=3"
    aff_dtype, s_a = torch.float32, 4


WORKLOADS = {"par": ParPseudoLabel, "par32": ParPseudoLabelF32}
try:
    from weclip_b200.bench_workloads import WORKLOADS as _MORE  # full-path workloads register themselves here
    WORKLOADS.update(_MORE)
except ImportError:
    pass
DEFAULT_WORKLOAD = "mscflip" if "mscflip" in WORKLOADS else "par"


def _config(wl_or_cls, world):
    cfg = {"workload": wl_or_cls.name, "images_per_step_per_gpu": getattr(wl_or_cls, "units", None),
           "parallelism": f"image-sharded x{world}", "l2": getattr(wl_or_cls, "l2_note", None)}
    if getattr(wl_or_cls, "cam_passes_per_image", None) is not None:
        cfg["cam_passes_per_image"] = wl_or_cls.cam_passes_per_image
    return cfg


def run_reference(args, rank, world):
    """Reference arm: the reference's own CPU implementation of the path (oracle port: the reference is Python and cannot
    travel) on all host cores of rank 0; a step = ONE dataset image of the same workload (a bounded sample of our arm's
    8-image step).  --steps / --warmup are honoured as given.  At N > 1 still ONE CPU process runs: the value is not
    scaled by N (stated in `note`)."""
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    cls = WORKLOADS[args.workload]
    wl = cls.cpu()
    for _ in range(args.warmup):
        wl.cpu_sample(1)
    vals = []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        v, sample = wl.cpu_sample(1)
        vals.append(v)
    dt = time.perf_counter() - t0
    v = statistics.mean(vals)
    cfg = _config(cls, world)        # the same keys and values as our arm's line
    unit = getattr(cls, "unit", "images/s")
    kind = getattr(cls, "cpu_kind", "port")
    line = {"impl": "reference", "metric": getattr(cls, "metric", METRIC), "value": round(v, 4), "unit": unit, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * dt / args.steps, 2),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg,
            "note": "one CPU process on rank 0 using all host cores; not multiplied by --gpus",
            "cpu_baseline": {"value": round(v, 4), "unit": unit, "cores": getattr(cls, "cpu_cores", None) or torch.get_num_threads(),
                             "kind": kind, "sample": sample},
            "e2e": {"value": round(v, 4), "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    from weclip_b200 import _lib
    _lib.require_device()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL prints its version banner on stdout at communicator creation: send it to stderr so stdout carries
        # nothing but the JSON line
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            os.dup2(saved, 1)
            os.close(saved)

    wl = WORKLOADS[args.workload](rank, dev)
    peaks = _peaks()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        wl.step_device()
        wl.step_e2e()
    barrier()

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    # --- device-resident timed region
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        wl.step_device(record=True)
    if hasattr(wl, "join"):
        wl.join()            # the step's passes run on the workload's own streams: order the timing stream after them
    e1.record()
    barrier()
    ms_dev = e0.elapsed_time(e1)
    # --- end-to-end timed region (host buffers, H2D + D2H inside)
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        wl.step_e2e()
    e1.record()
    barrier()
    ms_e2e = max(e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0))
    hist_total_e2e = int(wl.hist_h.sum()) if hasattr(wl, "hist_h") else None
    # --- the same step with dead CAM / PAR passes eliminated (cam_passes="live")
    ms_live = None
    if hasattr(wl, "ev_live"):
        for _ in range(2):
            wl.step_device_live()
        barrier()
        e0.record()
        for _ in range(args.steps):
            wl.step_device_live()
        wl.join()
        e1.record()
        barrier()
        ms_live = e0.elapsed_time(e1)
    # --- the same step with fp32 PAR affinities (the reference's PAR precision)
    ms_f32par = None
    if hasattr(wl, "ev") and hasattr(wl, "step_device_fp32_par") and getattr(wl, "cam_passes_per_image", 0):
        for _ in range(2):
            wl.step_device_fp32_par()
        barrier()
        e0.record()
        for _ in range(args.steps):
            wl.step_device_fp32_par()
        wl.join()
        e1.record()
        barrier()
        ms_f32par = e0.elapsed_time(e1)
    # --- one dataset image per call (the reference's batch_size = 1 loop), 512x512 and a VOC-like size mix
    ms_b1 = ms_b1v = None
    if hasattr(wl, "ev_b1"):
        for fn, tag in ((wl.step_b1, "b1"), (wl.step_b1_voc_sizes, "b1v")):
            for _ in range(3):
                fn()
            barrier()
            e0.record()
            for _ in range(args.steps):
                fn()
            wl.join()
            e1.record()
            barrier()
            if tag == "b1":
                ms_b1 = e0.elapsed_time(e1)
            else:
                ms_b1v = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    # --- kernel-class breakdown for the roofline object: separate pass with the C-side profiler (events around
    #     every launch on the launching stream); never overlaps the timed regions above
    if hasattr(wl, "profile"):
        wl.profile(max(1, min(args.steps, 3)))
        barrier()

    if dist is not None:
        t = torch.tensor([ms_dev, ms_e2e, ms_live or 0.0, ms_b1 or 0.0, ms_b1v or 0.0, ms_f32par or 0.0], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e, ms_live_r, ms_b1_r, ms_b1v_r, ms_f32par_r = t.tolist()
        ms_f32par = ms_f32par_r if ms_f32par is not None else None
        ms_live = ms_live_r if ms_live is not None else None
        ms_b1 = ms_b1_r if ms_b1 is not None else None
        ms_b1v = ms_b1v_r if ms_b1v is not None else None

    if rank == 0:
        units = wl.units * world * args.steps
        unit = getattr(wl, "unit", "images/s")
        line = {
            "metric": getattr(wl, "metric", METRIC), "value": round(units / (ms_dev * 1e-3), 2), "unit": unit, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_dev / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": getattr(wl, "dtype", "f32"),
            "data": "synthetic (seeded images / CAMs / random-init weights; no datasets or checkpoints offline)",
            "config": _config(wl, world),
            "clocks": clocks,
            "e2e": {"value": round(units / (ms_e2e * 1e-3), 2), "unit": unit,
                    "h2d_bytes_per_step": wl.h2d, "d2h_bytes_per_step": wl.d2h},
            "gpu_launches": int(wl.launches) * args.steps,
            "roofline": wl.roofline(peaks),
        }
        if hasattr(wl, "graphs"):
            line["config"]["cuda_graphs"] = bool(wl.graphs)
        if ms_live is not None:
            line["value_live_outputs"] = round(units / (ms_live * 1e-3), 2)
            line["value_live_outputs_note"] = ("same step, cam_passes='live': 1 CAM/PAR pass per dataset image instead of the "
                                               "reference call pattern's 6; identical histograms")
        if ms_f32par is not None:
            line["value_fp32_par_affinity"] = round(units / (ms_f32par * 1e-3), 2)
            line["value_fp32_par_affinity_note"] = ("same step with the PAR affinities stored in fp32 (the reference's PAR precision, "
                                                    "max error 2e-5); `value` stores them in fp16 (2.3e-3, labels >= 99.5 %)")
        if ms_b1 is not None:
            line["value_b1"] = round(units / (ms_b1 * 1e-3), 2)
            line["value_b1_voc_sizes"] = round(units / (ms_b1v * 1e-3), 2)
            line["value_b1_note"] = ("one dataset image per model call (the reference's DataLoader(batch_size=1)), reference call "
                                     "pattern; value_b1: the step's 512x512 images; value_b1_voc_sizes: 8 images of VOC-like sizes "
                                     f"{sorted(set(wl.voc_sizes))} (long side 500 -> 512, short sides mostly not multiples of 16); "
                                     f"CUDA graphs per bucket: {'on' if wl.graphs else 'off'}, graphs captured "
                                     f"{wl.ev_b1.graph_captures}, replays {wl.ev_b1.graph_replays}")
            line["config"]["cuda_graphs"] = bool(wl.graphs)
        if hist_total_e2e is not None:
            # the e2e region ends with the cross-rank histogram sum: every rank's step counts the same number of valid pixels
            line["e2e"]["hist_total_after_allreduce"] = hist_total_e2e
            line["e2e"]["collective"] = "all_reduce(SUM) of the 3 CxC int64 histograms inside the timed region" if world > 1 \
                else "none at world size 1"
        if world == 1 and hasattr(wl, "gpu_library_sample") and os.environ.get("WC_BENCH_GPU_LIB", "1") != "0":
            try:
                v, sample = wl.gpu_library_sample()
                line["gpu_library_baseline"] = {"value": round(v, 3), "unit": "images/s", "kind": "oracle loop on cuda "
                                                "(ATen/cuBLAS library calls)", "sample": sample}
            except Exception as exc:  # noqa: BLE001 - a baseline leg must never take the bench line down
                line["gpu_library_baseline"] = {"value": None, "error": repr(exc)[:200]}
        if world == 1:
            torch.set_num_threads(os.cpu_count() or 1)
            v, sample = wl.cpu_sample()
            line["cpu_baseline"] = {"value": round(v, 4), "unit": unit,
                                    "cores": getattr(wl, "cpu_cores", None) or torch.get_num_threads(),
                                    "kind": getattr(wl, "cpu_kind", "port"), "sample": sample}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
