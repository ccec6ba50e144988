#!/usr/bin/env python
"""Benchmark of the dense-stereo hot path (BASELINE.json metric: Gdisp/s = W*H*D / s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c5|c3|c4|c1] [--pairs P]

One step = one pass of census 9x7 -> Hamming cost volume (u16) -> SGM 8 directions -> WTA -> disparity over one
synthetic rectified pair (BASELINE.json configs[1]: 1920x1080, D=256; `--workload c5`: 1280x720, D=128).
  value  whole-job Gdisp/s with the pair already resident in HBM; the cost volume, the aggregated SGM volume and the
         disparity map are all written to HBM (what the reference's API returns as arrays).
  e2e    the same call through the C ABI with HOST operands: pinned host images in, host disparity map out, copies inside
         the timed region (the volumes stay in the device workspace). The caller's int32 map crosses PCIe either as it is or
         narrow (u8 / u16, widened by host threads): one ctx setting, same call, same array; both are timed and the line's
         e2e is the faster one on this box at this rank count (`d2h_wire` says which, `variants` holds both).
  roofline  the dominant kernel (SGM apply + fused WTA), timed per launch with CUDA events on the launching stream.
  cpu_baseline / --impl reference  the reference's own CPU code (oracle/_ref, compiled from the unmodified LibStevi
         headers) on the host cores, on a bounded row band of the same workload.
N > 1 (torchrun): independent pairs are sharded across ranks, no data-path collective (weak scaling).
The default (c2) line also carries `sustained` (>= 2 s of back-to-back pairs, clocks sampled throughout) and `secondary`: the
other configurations (c1, c3, c4, c5 incl. a streamed batch of distinct pairs, and the general SGM recurrence as a stage) with
fewer steps each. `--workload c5 --pairs 1024` streams the batch of BASELINE configs[4] and reports pairs/s.
Secondary workloads on their own: `--workload c3` = ZNCC 9x9 cost volume + WTA + refinement (configs[2]);
`--workload c4` = the coarse-to-fine driver on one 4096x3072 frame, split in row bands over the ranks and gathered with
one NCCL all_gather (configs[3], strong scaling).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# the library's host-side widening threads (OpenMP) must sleep between pairs, not spin on the cores that feed the copy engines
os.environ.setdefault("OMP_WAIT_POLICY", "PASSIVE")
os.environ.setdefault("GOMP_SPINCOUNT", "0")

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (H, W, D, h_radius, v_radius, seed, d_max)
    "c2": dict(H=1080, W=1920, D=256, hr=4, vr=3, seed=2, d_max=200,
               name="census9x7+SGM8+WTA 1920x1080 D=256 u16 costs (BASELINE configs[1])"),
    "c5": dict(H=720, W=1280, D=128, hr=4, vr=3, seed=1000, d_max=100,
               name="census9x7+SGM8+WTA 1280x720 D=128 u16 costs (BASELINE configs[4], one pair per step per GPU)"),
    "c3": dict(kind="zncc", H=1080, W=1920, D=192, hr=4, vr=4, seed=3, d_max=120,
               name="ZNCC9x9 cost volume + WTA + truncated volume + parabola refinement 1920x1080 D=192 f32 (BASELINE configs[2])"),
    "c1": dict(kind="zncc", H=480, W=640, D=64, hr=2, vr=2, seed=1, d_max=0,
               name="ZNCC5x5 cost volume + WTA (+ radius-1 refinement) 640x480 D=64 f32 (BASELINE configs[0], its dense form: SURVEY 3.5)"),
    "c4": dict(kind="hier", H=3072, W=4096, D=512, hr=3, vr=3, seed=4, d_max=0, depth=3, radius=2,
               name="hierarchical ZNCC7x7 depth 3, 4096x3072 D=512 f32, one frame split in row bands over the GPUs (BASELINE configs[3])"),
}
P1, P2, POUT, NDIR = 10.0, 120.0, 100.0, 8
# algorithmic bytes per disparity cell. CV: 2 written; SGM: 2 read + 4 written; WTA: 4 read.
BYTES_PER_CELL_PIPELINE = 12
BYTES_PER_CELL_SGM_APPLY = 6  # the dominant kernel: reads C (u16), writes S (f32); the WTA rides along for free


def read_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append([x.strip() for x in ln.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows if len(r) >= 7 for k in range(4) if r[3 + k].lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_reference_run(wl, steps, warmup, band_rows=None, budget_s=15.0):
    """The reference's own CPU implementation (oracle/_ref: unmodified LibStevi headers) on a row band of the workload."""
    import oracle as O
    from libstevi_b200 import synth
    chk, kind = (O.ref, "reference") if O.ref is not None else (O.port, "port")
    cores = os.cpu_count() or 1
    chk.set_num_threads(cores)
    H, W, D = wl["H"], wl["W"], wl["D"]
    L, R, _ = synth.stereo_pair(H, W, wl["seed"], d_bg=8, d_sq=32, d_max=wl["d_max"])
    if band_rows is None:
        # the reference SGM is O(D^2) per pixel and line (~0.3 s per 1920-pixel row at D=256 on 8 cores): size the band
        # so that all (warmup + steps) passes take about `budget_s` seconds of CPU work
        t_row = 0.31 * (D / 256.0) ** 2 * (W / 1920.0) * 8.0 / cores
        band_rows = int(budget_s / (max(steps + warmup, 1) * t_row))
        band_rows = max(4, min(band_rows, H))
    i0 = (H - band_rows) // 2
    Lb, Rb = np.ascontiguousarray(L[i0:i0 + band_rows]), np.ascontiguousarray(R[i0:i0 + band_rows])

    def step():
        cv = chk.cost_volume(O.CENSUS, Lb, Rb, wl["hr"], wl["vr"], D)
        sg = chk.sgm(cv, NDIR, O.COST, P1, P2, POUT)
        return chk.index_to_disp(chk.wta(sg, O.COST))

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / max(steps, 1)
    cells = band_rows * W * D
    return {"value": cells / dt / 1e9, "unit": "Gdisp/s", "cores": cores, "kind": kind, "band_rows": band_rows, "frame_rows": H,
            "full_frame_s_extrapolated": round(dt * H / band_rows, 1),
            "sample": f"rows {i0}..{i0 + band_rows} of the {W}x{H} pair ({band_rows}x{W}xD{D} cells per step), "
                      f"{steps} step(s), {dt:.2f} s/step"}, dt


def cpu_reference_float(wl, steps, warmup, budget_s=15.0):
    """The reference's CPU code for the float workloads on a bounded sample: c3 = a row band of the 1080p pair through
    cost volume + WTA + truncated volume + refinement; c4 = the whole hierarchical driver on a reduced frame."""
    import oracle as O
    from libstevi_b200 import synth
    chk, kind = (O.ref, "reference") if O.ref is not None else (O.port, "port")
    cores = os.cpu_count() or 1
    chk.set_num_threads(cores)
    n = max(steps + warmup, 1)
    if wl["kind"] == "zncc":
        H, W, D = wl["H"], wl["W"], wl["D"]
        L, R, _ = synth.stereo_pair(H, W, wl["seed"], d_bg=8, d_sq=32, d_max=wl["d_max"], dtype=np.float32)
        t_row = 0.035 * 8.0 / cores  # ~35 ms per 1920x192 row of 9x9 ZNCC on 8 cores (measured in the survey probe)
        rows = max(2 * wl["vr"] + 2, min(H, int(budget_s / (n * t_row))))
        i0 = (H - rows) // 2
        Lb, Rb = np.ascontiguousarray(L[i0:i0 + rows]), np.ascontiguousarray(R[i0:i0 + rows])

        def step():
            cv = chk.cost_volume(O.ZNCC, Lb, Rb, wl["hr"], wl["vr"], D)
            idx = chk.wta(cv, O.SCORE)
            return chk.refine(chk.truncated_cost_volume(cv, idx, wl["hr"], wl["vr"], 1), idx, O.PARABOLA)
        cells = rows * W * D
        sample = f"rows {i0}..{i0 + rows} of the {W}x{H} pair ({rows}x{W}xD{D} cells per step)"
    else:
        H, W, D = wl["H"] // 8, wl["W"] // 8, wl["D"] // 8
        L, R, _ = synth.stereo_pair(H, W, wl["seed"], d_bg=3, d_sq=12, dtype=np.float32)

        def step():
            return chk.hierarchical(O.ZNCC, wl["depth"], L, R, wl["hr"], wl["vr"], D, wl["radius"])
        cells = H * W * D
        sample = f"the same driver on a {W}x{H} D={D} frame (1/8 scale per axis; {cells} cells per step)"
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / max(steps, 1)
    return {"value": cells / dt / 1e9, "unit": "Gdisp/s", "cores": cores, "kind": kind,
            "sample": f"{sample}, {steps} step(s), {dt:.2f} s/step"}, dt




def timed_pair_stream(api, torch, call, outs, steps, warmup, device, barrier, flush, narrow_wire=True):
    """End-to-end throughput with len(outs) pairs in flight (api.HostPairStream: one ctx + stream per slot, deferred host
    completion): every step enqueues its own H2D copies from pinned memory, its kernels and its D2H copy; the copies of one
    pair overlap the kernels of another. Device time from the first enqueue to the last result landing in host memory
    (event pair on the idle default stream, recorded after a full synchronise and after the drain). No L2 flush is possible
    between overlapped steps; each step streams >= 1.6 GB of cost volume through the 126 MB L2."""
    hs = api.HostPairStream(len(outs), device, narrow_wire=narrow_wire)
    for _ in range(max(warmup, 3 * len(outs))):   # a ctx's workspace settles on its third call (grow, consolidate, steady)
        slot = hs.slot
        call(hs.acquire(), outs[slot])
    hs.drain()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    flush.zero_()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        slot = hs.slot
        call(hs.acquire(), outs[slot])
    hs.drain()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    hs.close()
    return ms


def main_float(args, wl, rank, world, local_rank, dist=None, steps=None, with_cpu=True):
    """c1 / c3 (ZNCC [+ refinement], independent pairs per rank) and c4 (coarse-to-fine on one frame split in row bands).
    Returns the JSON line as a dict on rank 0 (None elsewhere)."""
    steps = steps or args.steps
    H, W, D = wl["H"], wl["W"], wl["D"]
    cells = H * W * D
    hier = wl["kind"] == "hier"
    metric = "Gdisp/s (W*H*D/s) " + ("hierarchical ZNCC" if hier else "ZNCC cost volume+WTA+refinement")
    config = {"workload": wl["name"], "H": H, "W": W, "D": D, "window": f"{2 * wl['hr'] + 1}x{2 * wl['vr'] + 1}",
              "sharding": ("row bands with halos in global coordinates, one all_gather of the result bands (NCCL)" if hier
                           else "independent pairs per rank, no collective")}
    if hier:
        config.update(depth=wl["depth"], upscale_disp_radius=wl["radius"])
    if args.impl == "reference":
        if rank != 0:
            return None
        base, dt = cpu_reference_float(wl, args.steps, args.warmup, budget_s=120.0)
        config["cpu_sample"] = base["sample"]
        return {"impl": "reference", "metric": metric, "value": base["value"], "unit": "Gdisp/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
                "scaling": "strong" if hier else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config, "cpu_baseline": base, "gpu_launches": 0,
                "e2e": {"value": base["value"], "unit": "Gdisp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    import torch
    from libstevi_b200 import api, synth, sharding
    dev = torch.device("cuda", local_rank)
    warmup = max(args.warmup, 3)
    seed = wl["seed"] + (0 if hier else rank)
    L, R, gt = synth.stereo_pair(H, W, seed, d_bg=24 if hier else 8, d_sq=96 if hier else 32, d_max=wl["d_max"], dtype=np.float32)
    ctx = api.Context(local_rank, torch_stream=True)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    hl, hr_ = torch.from_numpy(L).pin_memory(), torch.from_numpy(R).pin_memory()
    if hier:
        bands = sharding.row_bands(H, world, 2)
        g0, g1 = bands[rank]
        plan = sharding.band_plan(H, wl["depth"], [wl["vr"]] * (wl["depth"] + 1), g0, g1)[-1]
        lo, hi = (0, H) if world == 1 else (plan.img_lo, plan.img_hi)
        dl, dr = torch.from_numpy(L[lo:hi]).to(dev), torch.from_numpy(R[lo:hi]).to(dev)
        res = {}

        def run(l, r):
            if world == 1:
                return api.hiearchicalTruncatedCostVolume(api.ZNCC, wl["depth"], l, r, wl["hr"], wl["vr"], D, wl["radius"], ctx=ctx)
            t, d = sharding.hierarchical_rows(api, api.ZNCC, wl["depth"], l, r, wl["hr"], wl["vr"], D, g0, g1, wl["radius"],
                                              band_row0=lo, full_height=H, ctx=ctx)
            return sharding.gather_rows(t, bands, dist), sharding.gather_rows(d, bands, dist)

        def step_device():
            res["tcv"], res["disp"] = run(dl, dr)

        def step_sharded():   # every rank keeps its band of the result: no collective at all
            res["tcv_band"], res["disp_band"] = sharding.hierarchical_rows(api, api.ZNCC, wl["depth"], dl, dr, wl["hr"], wl["vr"], D, g0, g1,
                                                                          wl["radius"], band_row0=lo, full_height=H, ctx=ctx)

        hdisp = torch.empty((H, W), dtype=torch.int32).pin_memory()

        def step_host():  # host frame in, host disparity map out (the band's rows travel H2D, the gathered map D2H)
            l, r = hl[lo:hi].to(dev, non_blocking=True), hr_[lo:hi].to(dev, non_blocking=True)
            _, d = run(l, r)
            hdisp.copy_(d, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        h2d, d2h = 2 * (hi - lo) * W * 4, H * W * 4
        dom_key, dom_bytes = "k_guided_cv", None
    else:
        dl, dr = torch.from_numpy(L).to(dev), torch.from_numpy(R).to(dev)
        cv = torch.empty((H, D, W), dtype=torch.float32, device=dev).permute(0, 2, 1)  # [H][D][W], cross_correlations.h:220
        idx = torch.empty((H, W), dtype=torch.int32, device=dev)
        tcv = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
        refined = torch.empty((H, W), dtype=torch.float32, device=dev)
        res = {"disp": idx}

        def step_device():
            api.localMatchingPipeline(api.ZNCC, dl, dr, wl["hr"], wl["vr"], D, api.PARABOLA, out=(cv, idx, tcv, refined), ctx=ctx)

        nl, nr = hl.numpy(), hr_.numpy()
        href = torch.empty((H, W), dtype=torch.float32).pin_memory()
        nref = href.numpy()

        def step_host():
            api.localMatchingPipeline(api.ZNCC, nl, nr, wl["hr"], wl["vr"], D, api.PARABOLA, out=nref, ctx=ctx)

        def call_stream(c, out):
            api.localMatchingPipeline(api.ZNCC, nl, nr, wl["hr"], wl["vr"], D, api.PARABOLA, out=out, ctx=c)
        stream_outs = [nref] + [torch.empty((H, W), dtype=torch.float32).pin_memory().numpy() for _ in range(args.inflight - 1)]
        h2d, d2h = 2 * H * W * 4, H * W * 4
        dom_key, dom_bytes = "k_cv_box", cells * 4

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step, steps, sampler=None):
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        if sampler:
            sampler.start()
        for e0, e1 in evs:
            flush.zero_()
            e0.record()
            step()
            e1.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms = sum(e0.elapsed_time(e1) for e0, e1 in evs)
        if dist is not None:
            ms = sharding.max_over_ranks(ms, dist, dev)
        return ms, clocks

    for _ in range(warmup):
        step_device()
    n0 = ctx.launches
    ms_dev, clocks = timed(step_device, steps, ClockSampler(local_rank) if rank == 0 else None)
    launches = ctx.launches - n0
    for _ in range(warmup):
        step_host()
    ms_host, _ = timed(step_host, steps)
    ms_serial, e2e_mode = ms_host, "one pair at a time, the call returns with the result in host memory"
    if not hier:
        ms_host = timed_pair_stream(api, torch, call_stream, stream_outs, steps, warmup, local_rank, barrier, flush)
        if dist is not None:
            ms_host = sharding.max_over_ranks(ms_host, dist, dev)
        e2e_mode = f"{args.inflight} pairs in flight (HostPairStream: one ctx+stream per slot, stevi_ctx_set_deferred_host)"
    ms_sharded = None
    if hier and world > 1:
        for _ in range(warmup):
            step_sharded()
        ms_sharded, _ = timed(step_sharded, steps)
    ctx.set_timing(True)
    per_kernel = {}
    for _ in range(max(3, min(steps, 10))):
        flush.zero_()
        step_device()
        for name, ms in ctx.timing_report():
            per_kernel.setdefault(name.split("<")[0].strip("()"), []).append(ms)
    ctx.set_timing(False)
    reps = max(3, min(steps, 10))
    kern_ms = {k: float(np.sum(v)) / reps for k, v in per_kernel.items()}   # per step, summed over the launches of a step
    torch.cuda.synchronize()
    d = res["disp"].cpu().numpy()
    inner = (slice(H // 3 + 40, 2 * H // 3 - 40), slice(W // 3 + 40, 2 * W // 3 - 40)) if hier else (slice(16, -16), slice(16, W - D))
    accuracy = float((d[inner] == gt[inner]).mean())
    line = None
    if rank == 0:
        peak, peak_src = read_peaks()
        units = cells * (1 if hier else world)   # c4: one frame in total (strong scaling); c3: one pair per rank
        line = {"metric": metric, "value": units * steps / (ms_dev * 1e-3) / 1e9, "unit": "Gdisp/s", "n_gpus": world,
                "steps": steps, "warmup": warmup, "ms_per_step": ms_dev / steps, "higher_is_better": True,
                "scaling": "strong" if hier else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "e2e": {"value": units * steps / (ms_host * 1e-3) / 1e9, "unit": "Gdisp/s", "h2d_bytes_per_step": int(h2d),
                        "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_host / steps, "mode": e2e_mode,
                        "serial_ms_per_step": ms_serial / steps},
                "gpu_launches": int(launches), "clocks": clocks, "kernels_ms_per_step": {k: round(v, 4) for k, v in kern_ms.items()},
                "ground_truth_agreement": accuracy}
        config["l2"] = "512 MB memset between timed steps (outside the event pairs)"
        if ms_sharded is not None:
            t_rows = 2 * wl["radius"] + 1
            line["result_sharded"] = {"ms_per_step": ms_sharded / steps, "value": units * steps / (ms_sharded * 1e-3) / 1e9, "unit": "Gdisp/s",
                                      "note": "same frame, same row bands, every rank keeps its band of (tcv, disp): no collective; the gathered "
                                              "form in `value` adds one NCCL all_gather of the bands",
                                      "gathered_bytes": int(H * W * (t_rows + 1) * 4),
                                      "gather_ms_per_step": (ms_dev - ms_sharded) / steps}
        dom = dom_key if dom_key in kern_ms else max(kern_ms, key=kern_ms.get)
        n_launch = len(per_kernel[dom]) / reps
        if dom_bytes:
            ach = dom_bytes / (kern_ms[dom] / n_launch * 1e-3) / 1e9
            traffic = None
            try:  # measured DRAM bytes of that kernel per launch, from the committed ncu --set full capture (taken on C3)
                with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                    traffic = json.load(f).get(dom) if (H, W, D) == (1080, 1920, 192) else None
            except Exception:
                traffic = None
            line["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": dom_bytes,
                                "ms_per_launch": kern_ms[dom] / n_launch}
            if dom == "k_cv_box":
                # the kernel also does extractSelectedIndex (the winner rides on it: the volume is never read back). Against the
                # bytes the two API stages move when run apart — SURVEY 8(d): CV written + WTA read, "fused, vs API-materialised
                # bytes" — the same launch time gives the second figure; `frac` stays the plain 4 bytes per cell.
                line["roofline"]["fused_stages"] = {
                    "stages": "unfoldBasedCostVolume (4 B/cell written) + extractSelectedIndex (4 B/cell read, fused away)",
                    "api_materialised_bytes_per_launch": 2 * dom_bytes, "achieved_vs_api_bytes": 2 * ach,
                    "frac_vs_api_bytes": 2 * ach / peak, "label": "fused, vs API-materialised bytes (may exceed 1)"}
        else:
            line["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": None, "peak": peak, "unit": "GB/s", "frac": None,
                                "traffic": None, "peak_source": peak_src, "ms_per_step": kern_ms[dom],
                                "note": "instruction-bound window sums over L2-resident images; no HBM roofline applies"}
        if world == 1 and with_cpu and not args.no_cpu_baseline:
            line["cpu_baseline"], _ = cpu_reference_float(wl, 1, 0, budget_s=15.0)
    if dist is not None:
        dist.barrier()
    ctx.close()
    del flush
    torch.cuda.empty_cache()
    return line


def main_census(args, wl, rank, world, local_rank, dist=None, steps=None, with_cpu=True):
    """c2 / c5: census -> Hamming cost volume -> SGM -> WTA, independent pairs per rank. Returns the line (rank 0)."""
    steps = steps or args.steps
    warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    H, W, D = wl["H"], wl["W"], wl["D"]
    cells = H * W * D
    config = {"workload": wl["name"], "H": H, "W": W, "D": D, "census_window": "9x7", "sgm_directions": NDIR,
              "P1": P1, "P2": P2, "Pout": POUT, "semantics": "compat (bit-exact with the reference CPU code)",
              "pairs_per_step_per_gpu": 1, "sharding": "independent pairs per rank, no collective"}

    if args.impl == "reference":
        if rank != 0:
            return None
        base, dt = cpu_reference_run(wl, steps, args.warmup, budget_s=150.0)
        # what this arm really ran: a row band of the frame per step (the reference's SGM is O(D^2) per pixel and line)
        config.update(cpu_band_rows=base["band_rows"], frame_rows=H, cpu_sample=base["sample"],
                      cpu_full_frame_s_extrapolated=round(dt * H / base["band_rows"], 1))
        line = {"impl": "reference", "metric": "Gdisp/s (W*H*D/s) census+SGM+WTA", "value": base["value"], "unit": "Gdisp/s",
                "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u16", "data": "synthetic",
                "config": config, "cpu_baseline": base,
                "e2e": {"value": base["value"], "unit": "Gdisp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        return line

    import torch
    from libstevi_b200 import api, synth
    dev = torch.device("cuda", local_rank)

    # each rank owns its own pair (different seed): weak scaling over independent pairs
    L, R, _ = synth.stereo_pair(H, W, wl["seed"] + rank, d_bg=8, d_sq=32, d_max=wl["d_max"])
    dl, dr = torch.from_numpy(L).to(dev), torch.from_numpy(R).to(dev)
    ctx = api.Context(local_rank, torch_stream=True)
    cv = torch.empty((H, D, W), dtype=torch.uint16, device=dev).permute(0, 2, 1)   # [H][D][W], cross_correlations.h:220
    sg = torch.empty((H, D, W), dtype=torch.float32, device=dev).permute(0, 2, 1)
    disp = torch.empty((H, W), dtype=torch.int32, device=dev)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step_device():
        api.censusSgmPipeline(dl, dr, wl["hr"], wl["vr"], D, P1, P2, POUT, NDIR, out=(cv, sg, disp), ctx=ctx)

    disp_only = torch.empty((H, W), dtype=torch.int32, device=dev)

    def step_device_disp():   # the same pipeline asked for the disparity map only: the aggregated volume is never stored
        api.censusSgmPipeline(dl, dr, wl["hr"], wl["vr"], D, P1, P2, POUT, NDIR, out=disp_only, ctx=ctx)

    hl, hr_ = torch.from_numpy(L).pin_memory(), torch.from_numpy(R).pin_memory()
    hdisp = torch.empty((H, W), dtype=torch.int32).pin_memory()
    nl, nr, nd = hl.numpy(), hr_.numpy(), hdisp.numpy()

    def step_host():
        api.censusSgmPipeline(nl, nr, wl["hr"], wl["vr"], D, P1, P2, POUT, NDIR, out=nd, ctx=ctx)

    def call_stream(c, out):
        api.censusSgmPipeline(nl, nr, wl["hr"], wl["vr"], D, P1, P2, POUT, NDIR, out=out, ctx=c)
    stream_outs = [nd] + [torch.empty((H, W), dtype=torch.int32).pin_memory().numpy() for _ in range(args.inflight - 1)]

    d2h_bytes = H * W * (1 if D <= 256 else 2 if D <= 65536 else 4)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step, steps, sampler=None):
        """Sum of per-step CUDA-event times; L2 flushed (512 MB memset) between steps, outside the event pairs."""
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        if sampler:
            sampler.start()
        for e0, e1 in evs:
            flush.zero_()
            e0.record()
            step()
            e1.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms = sum(e0.elapsed_time(e1) for e0, e1 in evs)
        if dist is not None:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, clocks

    for _ in range(warmup):
        step_device()
    n0 = ctx.launches
    ms_dev, clocks = timed(step_device, steps, ClockSampler(local_rank) if rank == 0 else None)
    launches = ctx.launches - n0

    for _ in range(warmup):
        step_host()
    for _ in range(warmup):
        step_device_disp()
    ms_disp, _ = timed(step_device_disp, steps)
    ms_serial, _ = timed(step_host, steps)
    ms_host = timed_pair_stream(api, torch, call_stream, stream_outs, steps, warmup, local_rank, barrier, flush)
    # the same loop with the int32 map itself on the wire, and with the caller taking the map as u8 / u16 (no widening)
    ms_wide = timed_pair_stream(api, torch, call_stream, stream_outs, steps, warmup, local_rank, barrier, flush, narrow_wire=False)
    ntype = np.uint8 if D <= 256 else np.uint16
    narrow_outs = [torch.empty((H, W), dtype=torch.uint8 if D <= 256 else torch.uint16).pin_memory().numpy() for _ in range(args.inflight)]
    ms_narrow = timed_pair_stream(api, torch, call_stream, narrow_outs, steps, warmup, local_rank, barrier, flush)
    narrow_same = bool(np.array_equal(narrow_outs[0].astype(np.int32), stream_outs[0]))
    if dist is not None:
        t = torch.tensor([ms_host, ms_wide, ms_narrow], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_host, ms_wide, ms_narrow = (float(x) for x in t.tolist())

    # sustained: back-to-back pairs for >= 2 s (no memset in between: every pair streams 3.2 GB through the 126 MB L2 anyway),
    # clocks sampled throughout (>= 20 samples at 50 ms)
    sustained = None
    if with_cpu:   # (the headline run only)
        n_sus = max(20, int(2.2e3 / max(ms_dev / steps, 1e-3)))
        sampler = ClockSampler(local_rank) if rank == 0 else None
        barrier()
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_sus):
            step_device()
        e1.record()
        barrier()
        sus_clocks = sampler.stop() if sampler else None
        ms_sus = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms_sus], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_sus = float(t.item())
        sustained = {"value": world * cells * n_sus / (ms_sus * 1e-3) / 1e9, "unit": "Gdisp/s", "pairs": n_sus, "seconds": ms_sus * 1e-3,
                     "ms_per_step": ms_sus / n_sus, "clocks": sus_clocks}

    # dominant kernel, per launch, CUDA events on the launching stream (second pass with the ctx's event hooks on)
    ctx.set_timing(True)
    per_kernel = {}
    for _ in range(max(3, min(steps, 10))):
        flush.zero_()
        step_device()
        for name, ms in ctx.timing_report():
            per_kernel.setdefault(name, []).append(ms)
    ctx.set_timing(False)
    kern_ms = {k: float(np.mean(v)) for k, v in per_kernel.items()}
    dom = max(kern_ms, key=kern_ms.get)
    peak, peak_src = read_peaks()
    dom_bytes = cells * BYTES_PER_CELL_SGM_APPLY if "sgm_apply" in dom else None
    traffic = None
    try:  # measured DRAM traffic of that kernel from the committed ncu --set full capture (C2 workload only)
        if (H, W, D) == (1080, 1920, 256):
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                traffic = json.load(f).get(dom.split("<")[0].strip("()"))
    except Exception:
        traffic = None

    # correctness guard inside the bench: device and host paths must agree bit for bit
    torch.cuda.synchronize()
    same = bool(np.array_equal(disp.cpu().numpy(), nd)) and bool(torch.equal(disp, disp_only))

    line = None
    if rank == 0:
        value = world * cells * steps / (ms_dev * 1e-3) / 1e9
        # the caller's int32 map can cross PCIe as it is or narrow (u8 / u16) and be widened by host threads: the same call, one ctx
        # setting (stevi_ctx_set_narrow_wire), the same array in the caller's hands. Which is faster depends on what is scarcer on
        # the box — PCIe bandwidth (many ranks copying: narrow wins) or host memory bandwidth / cores (one rank: the DMA engines
        # write the int32 map for free). Both are timed (max over ranks each); the line's e2e is the faster one and says which.
        narrow_pick = ms_host <= ms_wide
        ms_e2e = ms_host if narrow_pick else ms_wide
        d2h_e2e = int(d2h_bytes) if narrow_pick else int(nd.nbytes)
        e2e = world * cells * steps / (ms_e2e * 1e-3) / 1e9
        line = {"metric": "Gdisp/s (W*H*D/s) census+SGM+WTA", "value": value, "unit": "Gdisp/s", "n_gpus": world,
                "steps": steps, "warmup": warmup, "ms_per_step": ms_dev / steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u16", "data": "synthetic", "config": config,
                "e2e": {"value": e2e, "unit": "Gdisp/s", "h2d_bytes_per_step": int(nl.nbytes + nr.nbytes),
                        "d2h_bytes_per_step": d2h_e2e, "ms_per_step": ms_e2e / steps,
                        "d2h_wire": (f"disparity map as u{8 * d2h_bytes // (H * W)} over PCIe, widened into the caller's int32 array by "
                                     "host threads inside the timed region (stevi_ctx_set_narrow_wire on)" if narrow_pick else
                                     "the caller's int32 disparity map itself over PCIe (stevi_ctx_set_narrow_wire off)") +
                                    ": the faster of the two wire settings on this box at this rank count, both under variants",
                        "pcie_GBps_per_gpu": round((nl.nbytes + nr.nbytes + d2h_e2e) / (ms_e2e / steps * 1e-3) / 1e9, 2),
                        "variants": {
                            "narrow_map_on_the_wire_host_widening": {"ms_per_step": ms_host / steps, "value": world * cells * steps / (ms_host * 1e-3) / 1e9,
                                                                     "d2h_bytes_per_step": int(d2h_bytes)},
                            "int32_map_on_the_wire": {"ms_per_step": ms_wide / steps, "value": world * cells * steps / (ms_wide * 1e-3) / 1e9,
                                                      "d2h_bytes_per_step": int(nd.nbytes)},
                            "caller_takes_a_narrow_map": {"ms_per_step": ms_narrow / steps, "value": world * cells * steps / (ms_narrow * 1e-3) / 1e9,
                                                          "d2h_bytes_per_step": int(narrow_outs[0].nbytes), "dtype": str(np.dtype(ntype)),
                                                          "equal_to_the_int32_map": narrow_same}},
                        "mode": f"{args.inflight} pairs in flight (HostPairStream: one ctx+stream per slot, stevi_ctx_set_deferred_host)",
                        "serial_ms_per_step": ms_serial / steps},
                "gpu_launches": int(launches), "clocks": clocks,
                "kernels_ms": {k.split("<")[0].strip("()"): round(v, 4) for k, v in kern_ms.items()},
                "pipeline_algorithmic_GBps": cells * BYTES_PER_CELL_PIPELINE / (ms_dev / steps * 1e-3) / 1e9,
                "disparity_only": {"ms_per_step": ms_disp / steps, "value": world * cells * steps / (ms_disp * 1e-3) / 1e9,
                                   "unit": "Gdisp/s", "note": "same pipeline, caller asks for the disparity map only: the aggregated "
                                   "costs are evaluated in registers by the winner-take-all pass and not stored (what the e2e call runs); "
                                   "`value` is with the cost volume and the aggregated volume materialised as the reference's calls return them"},
                "sustained": sustained, "host_device_paths_agree": same}
        config["l2"] = "512 MB memset between timed steps (outside the event pairs); per-step working set 3.2 GB >> 126 MB L2"
        if dom_bytes:
            ach = dom_bytes / (kern_ms[dom] * 1e-3) / 1e9
            line["roofline"] = {"bound": "hbm", "kernel": dom.split("<")[0].strip("()"), "achieved": ach, "peak": peak,
                                "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                                "algorithmic_bytes_per_launch": dom_bytes, "ms_per_launch": kern_ms[dom]}
        else:
            line["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": None, "peak": peak, "unit": "GB/s", "frac": None,
                                "traffic": None, "peak_source": peak_src}
        if world == 1 and with_cpu and not args.no_cpu_baseline:
            line["cpu_baseline"], _ = cpu_reference_run(wl, 1, 0, budget_s=15.0)
            config.update(cpu_band_rows=line["cpu_baseline"]["band_rows"], frame_rows=H)
    if dist is not None:
        dist.barrier()
    ctx.close()
    del flush, cv, sg
    torch.cuda.empty_cache()
    return line if rank == 0 else None


def main_c5_batch(args, wl, rank, world, local_rank, dist=None, pairs=1024):
    """BASELINE configs[4] as the batch it is: `pairs` DISTINCT synthetic 1280x720 pairs (seeds 1000 ...), sharded over the ranks
    (pair k -> rank k mod N, sharding.shard_pairs), streamed from page-locked host memory through the C ABI with
    `--inflight` pairs in flight per GPU; every result lands in its own host array. The loop of
    test/benchmarks/benchmarkCostVolumeWorkflow.cpp:175-260 with the calls left in flight."""
    import torch
    from concurrent.futures import ThreadPoolExecutor
    from libstevi_b200 import api, synth, sharding
    dev = torch.device("cuda", local_rank)
    H, W, D = wl["H"], wl["W"], wl["D"]
    mine = sharding.shard_pairs(pairs, world, rank)
    Ls = torch.empty((len(mine), H, W), dtype=torch.uint8).pin_memory()
    Rs = torch.empty((len(mine), H, W), dtype=torch.uint8).pin_memory()
    odt = {"int32": torch.int32, "u8": torch.uint8, "u16": torch.uint16}[args.result_dtype]
    out = torch.empty((len(mine), H, W), dtype=odt).pin_memory()
    nL, nR, nO = Ls.numpy(), Rs.numpy(), out.numpy()

    def gen(k):
        nL[k], nR[k], _ = synth.stereo_pair(H, W, wl["seed"] + mine[k], d_bg=8, d_sq=32, d_max=wl["d_max"])
    with ThreadPoolExecutor(min(8, os.cpu_count() or 1)) as ex:
        list(ex.map(gen, range(len(mine))))
    hs = api.HostPairStream(args.inflight, local_rank, narrow_wire=not args.wide_wire)

    def run(n):
        for k in range(n):
            api.censusSgmPipeline(nL[k], nR[k], wl["hr"], wl["vr"], D, P1, P2, POUT, NDIR, out=nO[k], ctx=hs.acquire())
        hs.drain()
    run(min(len(mine), 3 * args.inflight))   # workspaces settle
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    run(len(mine))
    e1.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    launches = hs.launches
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # spot check: the first and the last pair of this rank against the device-resident call
    ctx = api.Context(local_rank, torch_stream=True)
    ok = True
    for k in (0, len(mine) - 1):
        d = api.censusSgmPipeline(torch.from_numpy(nL[k]).to(dev), torch.from_numpy(nR[k]).to(dev), wl["hr"], wl["vr"], D, P1, P2, POUT, NDIR, ctx=ctx)
        ok = ok and bool(np.array_equal(d.cpu().numpy(), nO[k].astype(np.int32)))
    ctx.close()
    hs.close()
    if rank != 0:
        return None
    per_pair_bytes = 2 * H * W + H * W * (4 if (args.wide_wire and args.result_dtype == "int32") else 2 if (D > 256 or args.result_dtype == "u16") else 1)
    return {"workload": wl["name"], "pairs": pairs, "n_gpus": world, "inflight_per_gpu": args.inflight, "seconds": ms * 1e-3,
            "pairs_per_s": pairs / (ms * 1e-3), "pairs_per_s_per_gpu": pairs / (ms * 1e-3) / world,
            "Gdisp_per_s": pairs * H * W * D / (ms * 1e-3) / 1e9, "ms_per_pair_per_gpu": ms / len(mine),
            "pcie_bytes_per_pair": per_pair_bytes, "pcie_GBps_per_gpu": round(per_pair_bytes * len(mine) / (ms * 1e-3) / 1e9, 2),
            "roofline_pairs_per_s_per_gpu_at_100pct": round(read_peaks()[0] * 1e9 / (12 * H * W * D), 1),
            "result_dtype": args.result_dtype, "wire": "int32 map" if (args.wide_wire and args.result_dtype == "int32") else "narrow map",
            "results_equal_device_resident_call": ok, "gpu_launches": int(launches)}


def sgm_general_stage(local_rank, steps=3):
    """The real SGM recurrence at C2 size (sgmCostVolume<8, Score> on a float volume of scores: the reference's own SGM benchmark
    case, benchmarkCrossCorrelationAlgorithms.cpp:282-303) - a stage figure, device resident, CUDA events."""
    import torch
    from libstevi_b200 import api
    dev = torch.device("cuda", local_rank)
    H, W, D = 1080, 1920, 256
    ctx = api.Context(local_rank, torch_stream=True)
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    v = (torch.rand((H, D, W), device=dev, generator=g) * 2 - 1).permute(0, 2, 1)
    fn = lambda: api.sgmCostVolume(v, 0.001, 0.01, Pout=100.0, nDirections=8, strategy=api.SCORE, ctx=ctx)
    for _ in range(2):
        fn()
    ts = []
    for _ in range(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ctx.set_timing(True); ctx.timing_report(); fn()
    kern = [[n.split("<")[0].strip("()"), round(ms, 4)] for n, ms in ctx.timing_report()]
    ctx.set_timing(False)
    ms = float(np.median(ts))
    alg = H * W * D * 8
    peak, _ = read_peaks()
    ctx.close()
    del v
    torch.cuda.empty_cache()
    return {"stage": "sgmCostVolume<8, Score> on a 1920x1080x256 f32 volume (general recurrence, sgm_sweep.cu)", "ms": ms,
            "Gcells_per_s": H * W * D / ms / 1e6, "algorithmic_bytes": alg, "algorithmic_GBps": alg / ms / 1e6,
            "frac_of_hbm_peak": alg / ms / 1e6 / peak, "kernels_ms": kern}


def pin_rank_to_cores(local_rank, world):
    """Each rank keeps to its own slice of the host cores (copy engines are fed by host threads; 8 ranks hopping over all
    cores of a virtual machine slow each other's copies down)."""
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = max(1, len(cores) // max(world, 1))
        mine = cores[local_rank * per:(local_rank + 1) * per] or cores
        os.sched_setaffinity(0, mine)
        os.environ.setdefault("STEVI_HOST_THREADS", str(max(1, min(4, len(mine)))))
        return len(mine)
    except Exception:
        return None


def copy_ceiling(rank, world, local_rank, dist, seconds=0.6):
    """What the box's host path gives when every rank copies at once (no kernels): pinned-memory H2D of one C2 image pair (4.1 MB)
    and D2H of one disparity map (8.3 MB as int32, 2.1 MB as u8), back to back on two streams. GB/s per GPU, min over ranks."""
    import torch
    dev = torch.device("cuda", local_rank)
    H, W = 1080, 1920
    out = {}
    for name, h2d_b, d2h_b in (("h2d_pair_only", 2 * H * W, 0), ("d2h_int32_map_only", 0, 4 * H * W), ("d2h_u8_map_only", 0, H * W),
                                ("pair_in_int32_map_out", 2 * H * W, 4 * H * W), ("pair_in_u8_map_out", 2 * H * W, H * W)):
        hin = torch.empty(max(h2d_b, 1), dtype=torch.uint8).pin_memory()
        din = torch.empty(max(h2d_b, 1), dtype=torch.uint8, device=dev)
        hout = torch.empty(max(d2h_b, 1), dtype=torch.uint8).pin_memory()
        dout = torch.empty(max(d2h_b, 1), dtype=torch.uint8, device=dev)
        s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

        def burst(n):
            for _ in range(n):
                if h2d_b:
                    with torch.cuda.stream(s1):
                        din.copy_(hin, non_blocking=True)
                if d2h_b:
                    with torch.cuda.stream(s2):
                        hout.copy_(dout, non_blocking=True)
        burst(20)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        n = 400
        t0 = time.perf_counter()
        burst(n)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if dist is not None:
            dist.barrier()
        gbps = n * (h2d_b + d2h_b) / dt / 1e9
        if dist is not None:
            t = torch.tensor([gbps], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            gbps = float(t.item())
        out[name] = {"GBps_per_gpu_min_over_ranks": round(gbps, 2), "ms_per_pair": round((h2d_b + d2h_b) / gbps / 1e6, 4)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--pairs", type=int, default=0, help="c5: stream this many distinct pairs (the batch of BASELINE configs[4])")
    ap.add_argument("--inflight", type=int, default=4, help="pairs in flight per GPU in the end-to-end measurements")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--result-dtype", default="int32", choices=["int32", "u8", "u16"], help="c5 batch: type of the host disparity maps")
    ap.add_argument("--wide-wire", action="store_true", help="c5 batch: send int32 maps over PCIe (no narrow wire)")
    ap.add_argument("--copy-ceiling", action="store_true", help="also measure what pinned copies alone reach with every rank copying")
    ap.add_argument("--no-pin", action="store_true", help="do not pin the rank to its slice of the host cores")
    ap.add_argument("--no-secondary", action="store_true", help="c2: leave the other configurations out of the line")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    is_float = wl.get("kind") in ("zncc", "hier")
    if args.impl == "reference":
        line = (main_float if is_float else main_census)(args, wl, rank, world, local_rank)
        if line is not None:
            print(json.dumps(line))
        return 0

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback (use --impl reference for the CPU arm)")
    cores = None if args.no_pin else pin_rank_to_cores(local_rank, world)
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if args.workload == "c5" and args.pairs > 0:
        line = main_c5_batch(args, wl, rank, world, local_rank, dist, args.pairs)
        if line is not None:
            line = {"metric": "pairs/s census+SGM D=128 1280x720 batch", "value": line["pairs_per_s"], "unit": "pairs/s", "n_gpus": world,
                    "higher_is_better": True, "scaling": "strong", "dtype": "u16", "data": "synthetic", "vs_baseline": None,
                    "config": {"workload": wl["name"], "pairs": args.pairs}, "batch": line}
    else:
        line = (main_float if is_float else main_census)(args, wl, rank, world, local_rank, dist)
    ceiling = copy_ceiling(rank, world, local_rank, dist) if args.copy_ceiling else None
    if line is not None:
        line["host_cores_per_rank"] = cores
        if ceiling:
            line["pinned_copy_ceiling"] = ceiling
    if args.workload == "c2" and not args.no_secondary and args.pairs == 0:
        # the other configurations, in the same line (fewer steps each; the driver only runs the default command)
        sec = {}
        k = max(3, min(args.steps, 8))
        for name in ("c1", "c3", "c4"):
            r = main_float(args, WORKLOADS[name], rank, world, local_rank, dist, steps=k, with_cpu=False)
            if r is not None:
                sec[name] = {q: r[q] for q in ("value", "unit", "ms_per_step", "scaling", "e2e", "roofline", "kernels_ms_per_step",
                                               "ground_truth_agreement", "gpu_launches", "result_sharded", "n_gpus") if q in r}
                sec[name]["workload"] = r["config"]["workload"]
        r = main_census(args, WORKLOADS["c5"], rank, world, local_rank, dist, steps=k, with_cpu=False)
        if r is not None:
            sec["c5"] = {q: r[q] for q in ("value", "unit", "ms_per_step", "e2e", "roofline", "gpu_launches") if q in r}
            sec["c5"]["workload"] = r["config"]["workload"]
        b = main_c5_batch(args, WORKLOADS["c5"], rank, world, local_rank, dist, pairs=128 * world)
        if b is not None:
            sec["c5"]["batch"] = b
        if world == 1:
            sec["sgm_general"] = sgm_general_stage(local_rank)
        if line is not None:
            line["secondary"] = sec
    if line is not None:
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
