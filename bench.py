#!/usr/bin/env python
"""bench.py -- denoise steps/s of the AnimateDiff denoise step (BASELINE.json metric) on N B200s.

    python bench.py --gpus N --steps K --warmup W            # our arm (sm_100a kernels via the C ABI)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle restatement)

Workload (config.workload): BASELINE.json configs[1] -- 16 frames 512x512 (latents [1,4,16,64,64]), CFG 7.5
(UNet batch 2), DDIM 25 steps, bf16 operands, random-init weights of the SD1.5 + motion-adapter v1.5-2
architecture, synthetic latents / text embeddings.  One "step" = one denoise step of one video: UNet on both CFG
halves + guidance combine + DDIM update (+ next step's model input).
N > 1: one process per GPU (torchrun).  --layout pairs: GPUs are grouped in CFG-split pairs (rank 2p = uncond half,
2p+1 = cond half, one NCCL all-gather of the noise predictions per step), pairs work on independent videos;
--layout replicas: every GPU runs its own video with both CFG halves (no data-path collective).
The reference arm runs the CPU restatement of the reference's loop body at the SAME shape, real steps, nothing
extrapolated.  Our arm's line also carries `parity` (rel-L2 of the noise prediction at the benchmarked shape against
the CPU oracle on identical weights) and, for pairs, `pair_latents_identical`.
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
# N > 1 layout when --layout auto: "pairs" = CFG-split GPU pairs (BASELINE configs[2], north star), "replicas" =
# independent single-GPU videos.  Both are measured (profiles/); the faster one is the default.
DEFAULT_LAYOUT = {2: "pairs", 4: "replicas", 8: "replicas"}


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
def _oracle_inputs(batch: int):
    import torch
    g = torch.Generator().manual_seed(42)
    lat = torch.randn(1, 4, FRAMES, LAT_H, LAT_W, generator=g)
    txt = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(43))
    return lat, (txt if batch == 2 else txt[1:])


def cpu_reference_steps(n_steps: int, warmup: int, unet=None):
    """The reference's CPU path (its diffusers loop body, restated in oracle/, kind 'port') on all host cores, at the
    workload's OWN shape: every step is one full CFG denoise step -- torch.cat([latents]*2) -> fp32 UNetMotionModel
    forward at batch 2 (16 frames, latents LAT_H x LAT_W) -> CFG combine -> DDIM update.  Nothing is extrapolated:
    `n_steps` real steps are timed after `warmup` real steps.  Returns (steps_per_s, ms list, info)."""
    import torch
    from oracle.unet_motion import make_oracle_unet
    from oracle.schedulers import OracleDDIM, cfg_combine
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    if unet is None:
        unet = make_oracle_unet(seed=0)
    lat, txt = _oracle_inputs(2)
    sch = OracleDDIM()
    sch.set_timesteps(NUM_STEPS)
    times = []
    with torch.no_grad():
        for i in range(warmup + n_steps):
            if i % NUM_STEPS == 0:
                x = lat.clone()
            t = int(sch.timesteps[i % NUM_STEPS])
            t0 = time.perf_counter()
            eps = unet(torch.cat([x] * 2), t, encoder_hidden_states=txt).sample
            x = sch.step(cfg_combine(eps, GUIDANCE), t, x)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt * 1e3)
    ms = sum(times) / len(times)
    info = {"cores": cores, "kind": "port",
            "sample": (f"{n_steps} full CFG denoise steps (fp32 UNet forward at batch 2, {FRAMES}f, latents {LAT_H}x{LAT_W}, "
                       f"+ CFG + DDIM) after {warmup} warm-up steps on {cores} threads; nothing extrapolated")}
    return 1e3 / ms, times, info


def cpu_half_forward_sample(unet, t: int):
    """Bounded cpu_baseline sample for our arm's line: ONE warm-up and ONE timed fp32 forward of the cond CFG half
    (B=1) at the workload's shape; a denoise step is two such halves (+ the negligible CFG/DDIM update, included).
    Returns (steps_per_s, info, eps of the cond half) -- the eps doubles as the parity reference of the bench shape."""
    import torch
    from oracle.schedulers import OracleDDIM
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    lat, txt = _oracle_inputs(1)
    sch = OracleDDIM()
    sch.set_timesteps(NUM_STEPS)
    with torch.no_grad():
        unet(lat, t, encoder_hidden_states=txt)                 # warm-up (page-in, oneDNN primitive caches)
        t0 = time.perf_counter()
        eps = unet(lat, t, encoder_hidden_states=txt).sample
        _ = sch.step(eps, t, lat)
        dt = time.perf_counter() - t0
    info = {"cores": cores, "kind": "port",
            "sample": (f"1 warm-up + 1 timed fp32 forward of one CFG half (B=1, {FRAMES}f, latents {LAT_H}x{LAT_W}) + DDIM "
                       f"update on {cores} threads; a step = 2 halves (value = 1 / (2 x sample time))")}
    return 1.0 / (2.0 * dt), info, eps


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    v, times, info = cpu_reference_steps(args.steps, args.warmup)
    ms = sum(times) / len(times)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": WORKLOAD.replace(", bf16", ", fp32 on the host CPU"), "device": "host CPU",
                       "threads": info["cores"], "latents": [1, 4, FRAMES, LAT_H, LAT_W], "unet_batch": 2},
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
    layout = args.layout
    if layout == "auto":
        layout = DEFAULT_LAYOUT.get(world, "pairs")
    if world == 1:
        layout = "single"
    pairs = layout == "pairs"
    pair_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        if pairs:
            if world % 2:
                raise SystemExit("CFG-split pairs need an even N")
            for p in range(world // 2):                      # every rank must create every group
                g = dist.new_group([2 * p, 2 * p + 1])
                if rank // 2 == p:
                    pair_group = g
    n_videos = world // 2 if pairs else world
    ok = False
    pipe = None
    try:
        pipe = AnimateDiffPipeline.from_pretrained("runwayml/stable-diffusion-v1-5", seed=0, with_vae=False)
        rc = _bench_body(args, pipe, pair_group, pairs, n_videos, world, rank, local)
        ok = rc == 0
    except BaseException:
        import traceback
        traceback.print_exc()
        if world == 1:
            raise
    if world > 1:
        from b200ad.dist_util import teardown_process_group
        return teardown_process_group(pipe, ok)
    return 0 if ok else 1


def _bench_body(args, pipe, pair_group, pairs, n_videos, world, rank, local):
    import torch
    import torch.distributed as dist
    from b200ad import lib
    from b200ad.ops import KernelProfiler, Ops
    from b200ad.schedulers import DDIMScheduler

    pipe.cfg_group = pair_group
    pipe.scheduler = DDIMScheduler.from_config(pipe.scheduler.config, beta_schedule="linear", steps_offset=1,
                                               clip_sample=False)
    pipe.scheduler.set_timesteps(NUM_STEPS)
    # CPU leg (rank 0, N = 1): the oracle UNet is built first and ITS (bf16-representable) random-init weights are
    # loaded into the product, so that the bounded CPU sample below is also the parity reference of the bench shape
    want_cpu = rank == 0 and world == 1 and not args.no_cpu_baseline
    oracle = None
    if want_cpu:
        from oracle.unet_motion import make_oracle_unet
        oracle = make_oracle_unet(seed=0)
        pipe.unet.load_state_dict(oracle.state_dict())
    pipe.unet.prepack()
    vid = (rank // 2 if pairs else rank) if world > 1 else 0
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
    clocks2 = ClockSampler(local)
    if rank == 0:
        clocks2.start()
    t0 = time.perf_counter()
    for j in range(k_e2e):
        e2e_step(j % NUM_STEPS)
    barrier()
    e2e_s = time.perf_counter() - t0
    clk2 = clocks2.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = n_videos * k_e2e / e2e_s
    h2d = lat_host.numel() * 4 + txt_host.numel() * 4
    d2h = out_host.numel() * 4

    # ---- CFG-split pair: are the two ranks' latents identical, and equal to the single-GPU step? --------
    pair_check = None
    if pairs:
        new_video()
        for i in range(3):
            plan.step(i)
        torch.cuda.synchronize()
        both = [torch.empty_like(plan.latents) for _ in range(2)]
        dist.all_gather(both, plan.latents, group=pair_group)
        identical = bool(torch.equal(both[0], both[1]))
        # the same three steps on ONE GPU with both CFG halves (eager, no collective), same inputs
        from b200ad.pipeline import DenoisePlan
        solo = DenoisePlan(pipe.unet, 2, FRAMES, LAT_H, LAT_W, use_graph=False, cfg_group=None)
        solo.load(lat_host.cuda(), text_dev, pipe.scheduler, GUIDANCE)
        for i in range(3):
            solo.step(i)
        torch.cuda.synchronize()
        d = (solo.latents - plan.latents).float()
        rel = float(d.norm() / solo.latents.float().norm())
        worst_rel = torch.tensor([rel], device="cuda")
        dist.all_reduce(worst_rel, op=dist.ReduceOp.MAX)
        ident_all = torch.tensor([1.0 if identical else 0.0], device="cuda")
        dist.all_reduce(ident_all, op=dist.ReduceOp.MIN)
        pair_check = {"pair_latents_identical": bool(ident_all.item() == 1.0), "steps": 3,
                      "pair_vs_single_gpu_rel_l2": float(worst_rel.item()),
                      "how": "3 graph steps on the pair; all_gather of latents inside each pair (bitwise compare); the "
                             "same 3 steps on one GPU with both CFG halves (eager); worst case over all ranks"}
        del solo

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
        # tensor-pipe fraction per attention head-dim family and HBM fraction of the memory-bound families
        for k, v in fam.items():
            if k.startswith("attention_"):
                v["frac_of_bf16_sustained"] = round(v["tflops"] / pk["bf16_tflops_sustained"], 4)
            elif k in ("group_norm", "layernorm", "temporal_attention", "cast", "cfg_step"):
                v["frac_of_hbm"] = round(v["gbs"] / pk["hbm_gbs"], 4)
        for v in fam.values():
            for k in ("ms", "tflops", "gbs"):
                v[k] = round(v[k], 4)

    # ---- CPU baseline + parity at the benchmarked shape (rank 0, N=1 only) ------------------------------
    cpu, parity = None, None
    if want_cpu:
        torch.cuda.synchronize()
        t_chk = int(pipe.scheduler._timesteps_host[0])
        v, info, eps_ref = cpu_half_forward_sample(oracle, t_chk)
        cpu = {"value": v, "unit": UNIT, **info}
        lat, txt = _oracle_inputs(1)
        eps = pipe.unet(lat.cuda(), t_chk, encoder_hidden_states=txt.cuda()).sample
        rel = float((eps.cpu() - eps_ref).norm() / eps_ref.norm())
        parity = {"rel_l2": rel, "gate": 1e-2, "ok": rel < 1e-2, "t": t_chk,
                  "what": f"noise prediction of the cond CFG half at the benchmarked shape (latents [1,4,{FRAMES},{LAT_H},"
                          f"{LAT_W}]) vs the fp32 CPU oracle on identical bf16-representable weights / latents / text"}
        if not parity["ok"]:
            print(f"bench.py: PARITY FAILURE at the benchmarked shape: rel-L2 {rel:.3e} > 1e-2", file=sys.stderr)

    if rank == 0:
        pk = _peaks()
        par = ("single GPU, both CFG halves" if world == 1 else
               f"{world // 2} CFG-split pair(s): NCCL all-gather of noise predictions per step, independent videos per pair"
               if pairs else f"{world} independent single-GPU replicas (both CFG halves per GPU), no data-path collective")
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": WORKLOAD, "latents": [1, 4, FRAMES, LAT_H, LAT_W], "unet_batch": 2,
                           "parallelism": par, "layout": "single" if world == 1 else ("pairs" if pairs else "replicas"),
                           "l2": "working set per step (2.6 GB bf16 weights + activations) >> 126 MB L2; no flush",
                           "cuda_graph": plan.graph is not None, "residual_stream": "fp32"},
                "videos_per_hour": value * 3600.0 / NUM_STEPS,
                "algorithmic_tflops_per_gpu": FLOPS_PER_STEP * value / 1e12 / world,
                "frac_of_bf16_sustained": FLOPS_PER_STEP * value / 1e12 / world / pk["bf16_tflops_sustained"],
                "clocks": clk,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "clocks": clk2,
                        "how": "per step: pinned-host latents+text -> device, text K/V, graph replay, latents -> host, sync"},
                "gpu_launches": int(launches_per_step * args.steps),
                "launches_per_step": int(launches_per_step),
                "roofline": roof, "families": fam, "cpu_baseline": cpu, "parity": parity}
        if pair_check is not None:
            line["pair_latents_identical"] = pair_check["pair_latents_identical"]
            line["pair_check"] = pair_check
        print(json.dumps(line), flush=True)
    if parity is not None and not parity["ok"]:
        return 3
    if pair_check is not None and not pair_check["pair_latents_identical"]:
        return 4
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
    ap.add_argument("--layout", default="auto", choices=["auto", "pairs", "replicas"],
                    help="N > 1: CFG-split pairs (one video per 2 GPUs) or independent single-GPU replicas")
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
