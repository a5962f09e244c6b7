#!/usr/bin/env python
"""bench.py -- denoise steps/s of the AnimateDiff denoise step (BASELINE.json metric) on N B200s.

    python bench.py --gpus N --steps K --warmup W            # our arm (sm_100a kernels via the C ABI)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle restatement)

Workload (config.workload): BASELINE.json configs[1] -- 16 frames 512x512 (latents [1,4,16,64,64]), CFG 7.5
(UNet batch 2), DDIM 25 steps, bf16 operands, random-init weights of the SD1.5 + motion-adapter v1.5-2
architecture, synthetic latents / text embeddings.  One "step" = one denoise step of one video: UNet on both CFG
halves + guidance combine + DDIM update (+ next step's model input).
N > 1: one process per GPU (torchrun); GPUs are grouped in CFG-split pairs (rank 2p = uncond half, 2p+1 = cond
half, one NCCL all-gather of the noise predictions per step); pairs work on independent videos.
Prints ONE JSON line on rank 0.
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
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "denoise steps/s (16f 512^2)"
UNIT = "steps/s"
WORKLOAD = "AnimateDiff SD1.5 UNet + motion-adapter v1.5 (random init), 16 frames 512x512, 25 DDIM steps, CFG 7.5, bf16"
FRAMES, LAT_H, LAT_W, NUM_STEPS, GUIDANCE = 16, 64, 64, 25, 7.5
FLOPS_PER_STEP = 35.498e12          # SURVEY.md Appendix C, algorithmic FLOPs of one CFG step at this shape
FLOPS_PER_STEP_C1 = 8.212e12        # the same at 16f 256^2 (BASELINE configs[0])


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"],
                "bf16_tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "source": "measured"}
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------- CPU reference path (oracle)
def cpu_reference_sample(n_samples: int, warmup: int, full_shape: bool):
    """Times the CPU restatement of the reference's diffusers path (oracle/, kind 'port') on all host cores.
    One sample = one fp32 UNetMotionModel forward of ONE CFG half (B=1, 16 frames) + the CFG combine and DDIM
    update; a denoise step is two such halves.  full_shape: latents 64x64 (the 512^2 workload itself);
    otherwise latents 32x32 (BASELINE configs[0], the reference's CPU-runnable case) with the step time
    extrapolated to 512^2 by the algorithmic FLOP ratio.  Returns (steps_per_s, ms list per step, info)."""
    import torch
    from oracle.unet_motion import make_oracle_unet
    from oracle.schedulers import OracleDDIM
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    unet = make_oracle_unet(seed=0, round_to=None)
    hw = LAT_H if full_shape else 32
    g = torch.Generator().manual_seed(42)
    lat = torch.randn(1, 4, FRAMES, hw, hw, generator=g)
    txt = torch.randn(1, 77, 768, generator=torch.Generator().manual_seed(43))
    sch = OracleDDIM()
    sch.set_timesteps(NUM_STEPS)
    scale = 1.0 if full_shape else FLOPS_PER_STEP / FLOPS_PER_STEP_C1
    times = []
    with torch.no_grad():
        for i in range(warmup + n_samples):
            t = int(sch.timesteps[i % NUM_STEPS])
            t0 = time.perf_counter()
            eps = unet(lat, t, encoder_hidden_states=txt).sample
            guided = eps + GUIDANCE * (eps - eps)
            _ = sch.step(guided, t, lat)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt * 2.0 * scale * 1e3)       # two CFG halves per denoise step
    ms = sum(times) / len(times)
    info = {"cores": cores, "kind": "port",
            "sample": (f"{n_samples} x one fp32 CFG-half UNet forward (B=1, 16f, latents {hw}x{hw}) + CFG + DDIM on "
                       f"{cores} threads; step = 2 halves" +
                       ("" if full_shape else f"; extrapolated to 512^2 by the FLOP ratio {scale:.3f}"))}
    return 1e3 / ms, times, info


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    full = (args.steps + args.warmup) <= 4
    v, times, info = cpu_reference_sample(args.steps, args.warmup, full_shape=full)
    ms = sum(times) / len(times)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "device": "host CPU", "threads": info["cores"]},
            "cpu_baseline": {"value": v, "unit": UNIT, **info},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------- our arm
def run_b200_arm(args):
    import torch
    import torch.distributed as dist
    import b200ad  # noqa: F401
    from b200ad import lib
    from b200ad.ops import KernelProfiler, Ops
    from b200ad.pipeline import AnimateDiffPipeline
    from b200ad.schedulers import DDIMScheduler

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the b200ad path has no CPU fallback)")
    torch.cuda.set_device(local)
    pair_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        if world % 2:
            raise SystemExit("N > 1 runs CFG-split pairs: N must be even")
        for p in range(world // 2):                      # every rank must create every group
            g = dist.new_group([2 * p, 2 * p + 1])
            if rank // 2 == p:
                pair_group = g
    n_videos = 1 if world == 1 else world // 2

    pipe = AnimateDiffPipeline.from_pretrained("runwayml/stable-diffusion-v1-5", seed=0, with_vae=False)
    pipe.cfg_group = pair_group
    pipe.scheduler = DDIMScheduler.from_config(pipe.scheduler.config, beta_schedule="linear", steps_offset=1,
                                               clip_sample=False)
    pipe.scheduler.set_timesteps(NUM_STEPS)
    pipe.unet.prepack()
    vid = rank // 2 if world > 1 else 0
    gen = torch.Generator().manual_seed(42 + vid)
    lat_host = torch.randn(1, 4, FRAMES, LAT_H, LAT_W, generator=gen).pin_memory()
    txt_host = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(43 + vid)).pin_memory()
    out_host = torch.empty_like(lat_host).pin_memory()
    plan = pipe._plan(2, FRAMES, LAT_H, LAT_W)
    text_dev = txt_host.cuda()
    my_text = text_dev[rank % 2][None] if plan.split else text_dev

    def new_video():
        plan.load(lat_host.cuda(non_blocking=True), my_text, pipe.scheduler, GUIDANCE)

    # launches per step (eager, counted by the library) -- also the first warm-up of every kernel
    new_video()
    plan.params.copy_(plan.table[0])
    n0 = lib.launch_count()
    plan._body()
    torch.cuda.synchronize()
    launches_per_step = lib.launch_count() - n0
    plan.capture()
    new_video()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run_steps(k, start=0):
        for j in range(k):
            i = (start + j) % NUM_STEPS
            if i == 0 and j > 0:
                new_video()
            plan.step(i)

    # ---- device-resident timing: W warm-up steps, then exactly K steps -------------------------------
    run_steps(args.warmup)
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run_steps(args.steps, start=args.warmup)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    clk = clocks.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_per_step = ms_total / args.steps
    value = n_videos * 1e3 / ms_per_step

    # ---- end to end through the public step API with HOST buffers --------------------------------------
    def e2e_step(i):
        lat = lat_host.cuda(non_blocking=True)
        txt = txt_host.cuda(non_blocking=True)
        plan.load(lat, txt[rank % 2][None] if plan.split else txt, pipe.scheduler, GUIDANCE)
        plan.step(i)
        out_host.copy_(plan.latents, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for i in range(min(3, args.warmup)):
        e2e_step(i)
    barrier()
    k_e2e = args.steps
    t0 = time.perf_counter()
    for j in range(k_e2e):
        e2e_step(j % NUM_STEPS)
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = n_videos * k_e2e / e2e_s
    h2d = lat_host.numel() * 4 + txt_host.numel() * 4
    d2h = out_host.numel() * 4

    # ---- per-kernel-family roofline: one eager step with CUDA events around every launch (rank 0) ------
    fam, roof = None, None
    if not args.no_profile:
        new_video()
        plan.params.copy_(plan.table[0])
        plan._body()                                       # warm caches after the e2e loop
        Ops.profiler = KernelProfiler()
        plan.params.copy_(plan.table[1])
        plan._body()
        fam = Ops.profiler.summary()
        Ops.profiler = None
        pk = _peaks()
        tens = {k: v for k, v in fam.items() if k in ("gemm", "conv3x3")}
        fl = sum(v["flops"] for v in tens.values())
        ms = sum(v["ms"] for v in tens.values())
        nl = sum(v["launches"] for v in tens.values())
        ach = fl / (ms * 1e-3) / 1e12
        # DRAM traffic of the same kernel from the committed ncu launch list (profiles/traffic.json, written by
        # tools/summarize_launches.py from `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum`): average
        # bytes per launch, like `achieved`; null when no capture is committed
        traffic, traffic_src = None, None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                tj = json.load(f)
            # the kernel is a template over its epilogue mode: gemm_tc_kernel, gemm_tc_kernel<0>, ... are one family
            gs = [v for k, v in tj.items() if k.split("<")[0] == "gemm_tc_kernel" and v.get("dram_bytes")]
            if gs:
                nl_t = sum(v["launches"] for v in gs)
                traffic = sum(v["dram_bytes"] for v in gs) / nl_t
                traffic_src = f"profiles/traffic.json ({nl_t} launches of one step under ncu)"
        except Exception:
            pass
        alg_bytes = sum(v["bytes"] for v in tens.values()) / max(nl, 1)
        roof = {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05 GEMM + implicit-GEMM conv3x3)",
                "achieved": ach, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                "frac": ach / pk["bf16_tflops_sustained"], "traffic": traffic, "traffic_unit": "bytes per launch",
                "traffic_source": traffic_src, "algorithmic_bytes_per_launch": alg_bytes,
                "peak_source": pk["source"] + " (sustained)",
                "launches_per_step": nl, "avg_launch_ms": ms / max(nl, 1),
                "share_of_step": ms / sum(v["ms"] for v in fam.values()),
                "how": "CUDA events around every launch of one eager step (no graph), algorithmic FLOPs 2MNK"}
        for v in fam.values():
            for k in ("ms", "tflops", "gbs"):
                v[k] = round(v[k], 4)

    # ---- CPU baseline (rank 0, N=1 only) --------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        torch.cuda.synchronize()
        v, _, info = cpu_reference_sample(1, 0, full_shape=True)
        cpu = {"value": v, "unit": UNIT, **info}

    if rank == 0:
        pk = _peaks()
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": WORKLOAD, "latents": [1, 4, FRAMES, LAT_H, LAT_W], "unet_batch": 2,
                           "parallelism": "single GPU, both CFG halves" if world == 1 else
                           f"{world // 2} CFG-split pair(s): NCCL all-gather of noise predictions per step, "
                           "independent videos per pair",
                           "l2": "working set per step (2.6 GB bf16 weights + activations) >> 126 MB L2; no flush",
                           "cuda_graph": plan.graph is not None, "residual_stream": "fp32"},
                "videos_per_hour": value * 3600.0 / NUM_STEPS,
                "algorithmic_tflops_per_gpu": FLOPS_PER_STEP * value / 1e12 / world,
                "frac_of_bf16_sustained": FLOPS_PER_STEP * value / 1e12 / world / pk["bf16_tflops_sustained"],
                "clocks": clk,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "how": "per step: pinned-host latents+text -> device, text K/V, graph replay, latents -> host, sync"},
                "gpu_launches": int(launches_per_step * args.steps),
                "launches_per_step": int(launches_per_step),
                "roofline": roof, "families": fam, "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
    if world > 1:
        # Tear-down: the captured CUDA graphs hold the pair communicator (the per-step all-gather is inside the graph);
        # destroying the process group with them alive hung at exit (first 2-GPU run of this bench: JSON line
        # printed, torchrun never returned).  Drop the graphs, drain the device, meet once more, then leave without
        # NCCL's finalizers.
        for pl in pipe._plans.values():
            pl.graph = None
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)
    return 0


def main():
    global FRAMES, LAT_H, LAT_W, NUM_STEPS, GUIDANCE, FLOPS_PER_STEP, WORKLOAD, METRIC
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=25)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    # other BASELINE configs (not the default bench line): --frames 32 --latent 96 = configs[4] (32f 768^2);
    # --num-steps 15 --guidance 9.0 with --gpus 2 = configs[2]
    ap.add_argument("--frames", type=int, default=FRAMES)
    ap.add_argument("--latent", type=int, default=LAT_H)
    ap.add_argument("--num-steps", type=int, default=NUM_STEPS)
    ap.add_argument("--guidance", type=float, default=GUIDANCE)
    args = ap.parse_args()
    if (args.frames, args.latent, args.num_steps, args.guidance) != (FRAMES, LAT_H, NUM_STEPS, GUIDANCE):
        from tools.flops import unet_flops
        FRAMES, LAT_H, LAT_W, NUM_STEPS, GUIDANCE = args.frames, args.latent, args.latent, args.num_steps, args.guidance
        FLOPS_PER_STEP = unet_flops(2, FRAMES, LAT_H)["total"]
        WORKLOAD = (f"AnimateDiff SD1.5 UNet + motion-adapter v1.5 (random init), {FRAMES} frames {8 * LAT_H}x{8 * LAT_W}, "
                    f"{NUM_STEPS} DDIM steps, CFG {GUIDANCE}, bf16")
        METRIC = f"denoise steps/s ({FRAMES}f {8 * LAT_H}^2)"
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
