"""The reference's 78-run grid-search sweep, data-parallel over CFG-split GPU pairs (BASELINE
configs[3]).

The reference runs the sweep sequentially in one process
(/root/reference/experiments/05_grid_search_ablation.py:195-334): per test video a CFG ablation (6 guidance
scales at 25 steps), a steps ablation (6 step counts at CFG 7.5) and a prompt ablation (baseline / enhanced), an
experiment being skipped when ``<out>/<experiment_id>/config.json`` already exists (``05:212-216``) -- which is
also why ``<video>_cfg7.5_steps25`` is generated once although two phases name it: 6 x 13 = 78 runs, 2130 denoise
steps (``/root/reference/outputs/05_grid_search/manifest.json``).

Here the runs are independent work items: they are assigned to workers (a worker = one GPU, or one CFG-split
pair of GPUs) longest-processing-time-first over their step counts, every worker walks its own list, and the
reference's resume marker keeps the protocol restartable.  There is no collective between workers; inside a
pair the only exchange is the per-step all-gather of the noise predictions (pipeline.DenoisePlan).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        -m b200ad.sweep --output outputs/05_grid_search
"""
from __future__ import annotations

import argparse
import json
import os
import time
from dataclasses import asdict, dataclass
from pathlib import Path
from typing import Dict, List, Optional, Sequence

CFG_VALUES = [5.0, 6.0, 7.0, 7.5, 8.0, 9.0]            # 05:40
STEPS_VALUES = [15, 20, 25, 30, 40, 50]               # 05:41
DEFAULT_CFG, DEFAULT_STEPS = 7.5, 25                   # 05:44-45
NUM_FRAMES, HEIGHT, WIDTH, SEED = 16, 512, 512, 42     # 05:48-51
_VIDEOS_JSON = os.path.join(os.path.dirname(os.path.abspath(__file__)), "sweep_videos.json")


@dataclass
class ExperimentConfig:
    """Same fields, order and defaults as the reference's dataclass (05:100-113): config.json is interchangeable."""
    experiment_id: str
    video_name: str
    prompt: str
    negative_prompt: str
    guidance_scale: float
    num_inference_steps: int
    phase: str
    seed: int = SEED
    num_frames: int = NUM_FRAMES
    height: int = HEIGHT
    width: int = WIDTH


def load_test_videos(path: Optional[str] = None) -> Dict[str, Dict[str, str]]:
    """The six test videos with baseline / enhanced prompts (data of 05:57-94), in the reference's order."""
    with open(path or _VIDEOS_JSON) as f:
        return json.load(f)


def sweep_runs(videos: Optional[Dict[str, Dict[str, str]]] = None, phase: str = "all",
               video_filter: Optional[str] = None) -> List[ExperimentConfig]:
    """Every experiment the reference's run_grid_search generates, in its order, de-duplicated by experiment_id
    the way its resume marker de-duplicates them (05:195-334)."""
    videos = videos if videos is not None else load_test_videos()
    if video_filter:
        videos = {k: v for k, v in videos.items() if video_filter in k}
    runs: List[ExperimentConfig] = []
    seen = set()

    def add(eid, name, prompt, neg, g, n, ph):
        if eid not in seen:
            seen.add(eid)
            runs.append(ExperimentConfig(eid, name, prompt, neg, g, n, ph))

    for name, v in videos.items():
        if phase in ("all", "cfg"):
            for g in CFG_VALUES:
                add(f"{name}_cfg{g:.1f}_steps{DEFAULT_STEPS}", name, v["prompt_baseline"], v["negative_baseline"], g,
                    DEFAULT_STEPS, "cfg_ablation")
        if phase in ("all", "steps"):
            for n in STEPS_VALUES:
                add(f"{name}_cfg{DEFAULT_CFG:.1f}_steps{n}", name, v["prompt_baseline"], v["negative_baseline"],
                    DEFAULT_CFG, n, "steps_ablation")
        if phase in ("all", "prompt"):
            for variant in ("baseline", "enhanced"):
                add(f"{name}_cfg{DEFAULT_CFG:.1f}_steps{DEFAULT_STEPS}_prompt_{variant}", name,
                    v[f"prompt_{variant}"], v[f"negative_{variant}"], DEFAULT_CFG, DEFAULT_STEPS, "prompt_ablation")
    return runs


def assign_lpt(runs: Sequence[ExperimentConfig], n_workers: int) -> List[List[ExperimentConfig]]:
    """Longest-processing-time-first over the step counts (a run costs num_inference_steps UNet evaluations):
    sort by steps descending (stable: ties keep the reference order), always give the next run to the least
    loaded worker (lowest index on ties).  Deterministic, so every rank computes the same assignment."""
    loads = [0] * n_workers
    out: List[List[ExperimentConfig]] = [[] for _ in range(n_workers)]
    for r in sorted(runs, key=lambda r: -r.num_inference_steps):
        w = min(range(n_workers), key=lambda i: (loads[i], i))
        out[w].append(r)
        loads[w] += r.num_inference_steps
    return out


def is_done(run: ExperimentConfig, output_dir: Path) -> bool:
    return (Path(output_dir) / run.experiment_id / "config.json").exists()      # the reference's marker, 05:212


def save_experiment(frames, run: ExperimentConfig, output_dir: Path) -> Path:
    """frames/frame_%04d.png, <id>.gif, then config.json LAST (it is the completion marker) -- 05:172-190.
    `frames`: list of PIL images (the reference's synchronous form; the sweep itself uses FrameExporter)."""
    exp_dir = Path(output_dir) / run.experiment_id
    frames_dir = exp_dir / "frames"
    frames_dir.mkdir(parents=True, exist_ok=True)
    if frames is not None:
        for i, fr in enumerate(frames):
            fr.save(frames_dir / f"frame_{i:04d}.png")
        frames[0].save(os.fspath(exp_dir / f"{run.experiment_id}.gif"), save_all=True, append_images=frames[1:],
                       optimize=False, duration=100, loop=0)
    _write_marker(run, exp_dir)
    return exp_dir


class FrameExporter:
    """Frame export overlapped with the next video.  The reference saves 16 PNGs and a GIF
    synchronously between videos (05:172-190); at ~1 s of GPU time per video that stall would be a large share of the
    sweep.  Here the writer rank hands the uint8 frames (device tensor [F,H,W,3]) over: they are copied to pinned host
    memory on a side stream, PNG / GIF encoding runs on a small thread pool (PIL releases the GIL while it compresses),
    and config.json -- the completion marker -- is written only after every file of the run is on disk."""

    def __init__(self, output_dir: Path, threads: int = 4, synchronous: bool = False):
        import concurrent.futures as cf
        import threading
        self.out = Path(output_dir)
        self.sync = synchronous
        # >= 3 threads: up to two per-run tasks wait on their frame tasks, which need a free thread of the same pool
        self.pool = None if synchronous else cf.ThreadPoolExecutor(max_workers=max(3, threads), thread_name_prefix="export")
        self.pending = []
        self.lock = threading.Lock()
        self.encode_seconds = 0.0        # summed over threads
        self.stall_seconds = 0.0         # time the generating thread spent waiting for the exporter
        self.copy_stream = None

    def _timed(self, fn, *a):
        t0 = time.perf_counter()
        fn(*a)
        with self.lock:
            self.encode_seconds += time.perf_counter() - t0

    def submit(self, frames_u8, run: "ExperimentConfig"):
        """frames_u8: uint8 [F,H,W,3] CUDA tensor (or None for marker-only runs)."""
        import torch
        exp_dir = self.out / run.experiment_id
        (exp_dir / "frames").mkdir(parents=True, exist_ok=True)
        if frames_u8 is None:
            _write_marker(run, exp_dir)
            return
        t0 = time.perf_counter()
        if self.sync:
            arr = frames_u8.cpu().numpy()
            self._timed(_encode_run, arr, run, exp_dir)
            self.stall_seconds += time.perf_counter() - t0
            return
        if self.copy_stream is None:
            self.copy_stream = torch.cuda.Stream()
        host = torch.empty(frames_u8.shape, dtype=torch.uint8, pin_memory=True)
        self.copy_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.copy_stream):
            host.copy_(frames_u8, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        frames_u8.record_stream(self.copy_stream)
        # bound the backlog: at most two videos in flight behind the generator
        while len(self.pending) >= 2:
            self.pending.pop(0).result()
        self.pending.append(self.pool.submit(self._export_async, host, ev, run, exp_dir))
        self.stall_seconds += time.perf_counter() - t0

    def _export_async(self, host, ev, run, exp_dir):
        ev.synchronize()
        arr = host.numpy()
        # per frame, on the pool's other threads: the PNG, and the palette quantisation of the GIF frame (PIL's GIF
        # writer spends ~90 % of its time there; pre-quantised "P" frames give a byte-identical file); then the GIF is
        # assembled here and the marker written after every file of the run is on disk
        pngs = [self.pool.submit(self._timed, _encode_png, arr[i], exp_dir / "frames" / f"frame_{i:04d}.png")
                for i in range(arr.shape[0])]
        quant = [self.pool.submit(self._timed_ret, _quantize_frame, arr[i]) for i in range(arr.shape[0])]
        frames = [q.result() for q in quant]
        self._timed(_save_gif, frames, exp_dir / f"{run.experiment_id}.gif")
        for f in pngs:
            f.result()
        _write_marker(run, exp_dir)

    def _timed_ret(self, fn, *a):
        t0 = time.perf_counter()
        r = fn(*a)
        with self.lock:
            self.encode_seconds += time.perf_counter() - t0
        return r

    def finish(self):
        t0 = time.perf_counter()
        for f in self.pending:
            f.result()
        self.pending.clear()
        if self.pool is not None:
            self.pool.shutdown(wait=True)
        self.stall_seconds += time.perf_counter() - t0


def _encode_png(frame, path):
    from PIL import Image
    Image.fromarray(frame).save(path)


def _quantize_frame(frame):
    """RGB -> 256-colour adaptive palette, exactly the conversion PIL's GIF writer applies to an RGB frame."""
    from PIL import Image
    return Image.fromarray(frame).convert("P", palette=Image.Palette.ADAPTIVE)


def _save_gif(frames, path):
    frames[0].save(os.fspath(path), save_all=True, append_images=frames[1:], optimize=False, duration=100, loop=0)


def _encode_gif(arr, path):
    _save_gif([_quantize_frame(f) for f in arr], path)          # same bytes as export_to_gif(frames, path) (05:181-182)


def _encode_run(arr, run, exp_dir):
    for i in range(arr.shape[0]):
        _encode_png(arr[i], exp_dir / "frames" / f"frame_{i:04d}.png")
    _encode_gif(arr, exp_dir / f"{run.experiment_id}.gif")
    _write_marker(run, exp_dir)


def _write_marker(run: "ExperimentConfig", exp_dir: Path):
    tmp = exp_dir / "config.json.tmp"
    with open(tmp, "w") as f:
        json.dump(asdict(run), f, indent=2)
    os.replace(tmp, exp_dir / "config.json")


def generate_manifest(output_dir: Path) -> dict:
    """manifest.json in the reference's format (05:341-375)."""
    manifest = {"grid_params": {"cfg_values": CFG_VALUES, "steps_values": STEPS_VALUES, "default_cfg": DEFAULT_CFG,
                                "default_steps": DEFAULT_STEPS}, "experiments": []}
    for exp_dir in sorted(Path(output_dir).iterdir()):
        cp = exp_dir / "config.json"
        if cp.exists():
            with open(cp) as f:
                c = json.load(f)
            manifest["experiments"].append({"experiment_id": c["experiment_id"], "video_name": c["video_name"],
                                            "cfg": c["guidance_scale"], "steps": c["num_inference_steps"],
                                            "phase": c["phase"]})
    with open(Path(output_dir) / "manifest.json", "w") as f:
        json.dump(manifest, f, indent=2)
    return manifest


def worker_layout(world: int, pair: bool):
    """(n_workers, ranks per worker).  pair=True: GPUs 2p, 2p+1 form worker p (CFG halves split); else one
    worker per GPU (no exchange at all; each GPU evaluates both halves)."""
    if pair and world >= 2 and world % 2 == 0:
        return world // 2, 2
    return world, 1


def run_worker(pipe, my_runs: Sequence[ExperimentConfig], output_dir: Path, writer: bool, decode: bool = True,
               log=print, sync_export: bool = False, export_threads: int = 4, metrics_dir: Optional[Path] = None) -> dict:
    """Walk one worker's list the way the reference's loops do (generate -> save -> marker).  In a CFG pair both
    ranks run the denoise loop (they stay in lock-step through the per-step all-gather); only `writer` decodes the
    latents (b200ad VAE -> uint8 frames on the device) and exports, overlapped with the next video (FrameExporter)."""
    import torch
    exporter = FrameExporter(output_dir, threads=export_threads, synchronous=sync_export) if writer else None
    done_steps, done_runs, t0 = 0, 0, time.perf_counter()
    t_denoise = t_decode = t_measure = 0.0
    lp = None
    if writer and decode and metrics_dir is not None:
        from .lpips import LpipsAlex
        from .metrics import measure_video
        lp = LpipsAlex.from_local_or_random().cuda()
        Path(metrics_dir).mkdir(parents=True, exist_ok=True)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for run in my_runs:
        if is_done(run, output_dir):
            log(f"  Skipping {run.experiment_id} (already exists)")
            continue
        generator = torch.Generator().manual_seed(run.seed)          # 05:156 torch.manual_seed(seed): CPU generator
        ev[0].record()
        out = pipe(prompt=run.prompt, negative_prompt=run.negative_prompt, num_frames=run.num_frames,
                   guidance_scale=run.guidance_scale, num_inference_steps=run.num_inference_steps,
                   height=run.height, width=run.width, generator=generator, output_type="latent")
        ev[1].record()
        frames = None
        if writer and decode:
            frames = pipe.vae.decode_video_u8(out.frames.float())   # [F, H, W, 3] uint8, still on the device
        ev[2].record()
        if writer:
            exporter.submit(frames, run)
        if lp is not None and frames is not None:
            # the reference's experiments/06 on the frames while they are still on the device (row f4)
            tm = time.perf_counter()
            m = measure_video(frames, lpips_fn=lp)
            rec = {"video_name": run.video_name, "experiment_id": run.experiment_id, "num_frames": m.pop("num_frames"),
                   "guidance_scale": run.guidance_scale, "num_inference_steps": run.num_inference_steps,
                   "phase": run.phase, **m}
            with open(Path(metrics_dir) / f"{run.experiment_id}_metrics.json", "w", encoding="utf-8") as f:
                json.dump(rec, f, indent=2)
            t_measure += time.perf_counter() - tm
        ev[2].synchronize()
        t_denoise += ev[0].elapsed_time(ev[1]) * 1e-3
        t_decode += ev[1].elapsed_time(ev[2]) * 1e-3
        done_steps += run.num_inference_steps
        done_runs += 1
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t0
    if exporter is not None:
        exporter.finish()
    return {"runs": done_runs, "steps": done_steps, "seconds": time.perf_counter() - t0, "generate_seconds": t_gen,
            "denoise_seconds": t_denoise, "decode_seconds": t_decode, "measure_seconds": t_measure,
            "export_encode_seconds": 0.0 if exporter is None else exporter.encode_seconds,
            "export_stall_seconds": 0.0 if exporter is None else exporter.stall_seconds}


def main(argv=None) -> int:
    import torch
    import torch.distributed as dist
    from .pipeline import AnimateDiffPipeline, MotionAdapter
    from .schedulers import DDIMScheduler

    ap = argparse.ArgumentParser(description="78-run grid-search sweep, data-parallel over B200s")
    ap.add_argument("--phase", default="all", choices=["all", "cfg", "steps", "prompt"])
    ap.add_argument("--video", default=None)
    ap.add_argument("--output", default="outputs/05_grid_search")
    ap.add_argument("--no-pairs", action="store_true", help="one worker per GPU instead of CFG-split pairs")
    ap.add_argument("--latent-only", action="store_true", help="skip VAE decode / frame export (markers only)")
    ap.add_argument("--limit", type=int, default=0, help="only the first LIMIT runs (smoke runs)")
    ap.add_argument("--sync-export", action="store_true",
                    help="encode PNG/GIF synchronously between videos, as the reference does (05:172-190)")
    ap.add_argument("--measure", default=None, metavar="DIR",
                    help="also write the reference's experiments/06 metrics (<id>_metrics.json) for every video into DIR")
    ap.add_argument("--export-threads", type=int, default=0,
                    help="encoder threads per writer rank (0: host cores / ranks on this node, at least 3)")
    ap.add_argument("--height", type=int, default=HEIGHT)
    ap.add_argument("--width", type=int, default=WIDTH)
    a = ap.parse_args(argv)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    group = None
    n_workers, per = worker_layout(world, not a.no_pairs)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        if per == 2:
            for p in range(n_workers):                    # every rank creates every group
                g = dist.new_group([2 * p, 2 * p + 1])
                if rank // 2 == p:
                    group = g
    worker = rank // per
    out_dir = Path(a.output)
    if rank == 0:
        out_dir.mkdir(parents=True, exist_ok=True)
    if world > 1:
        dist.barrier()

    runs = sweep_runs(phase=a.phase, video_filter=a.video)
    if a.limit:
        runs = runs[:a.limit]
    for r in runs:
        r.height, r.width = a.height, a.width
    # the pending set is fixed once, before anyone writes a marker, so all ranks agree on the assignment
    pending = [r for r in runs if not is_done(r, out_dir)]
    if world > 1:
        dist.barrier()
    mine = assign_lpt(pending, n_workers)[worker]

    adapter = MotionAdapter.from_pretrained("guoyww/animatediff-motion-adapter-v1-5-2")      # 05:124
    pipe = AnimateDiffPipeline.from_pretrained("runwayml/stable-diffusion-v1-5", motion_adapter=adapter,
                                               with_vae=not a.latent_only)                    # 05:130
    pipe.scheduler = DDIMScheduler.from_config(pipe.scheduler.config, beta_schedule="linear", steps_offset=1,
                                               clip_sample=False)                             # 05:136-141
    pipe.cfg_group = group
    if a.export_threads <= 0:
        a.export_threads = max(3, (os.cpu_count() or 8) // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", world))))
    ok = False
    try:
        stats = run_worker(pipe, mine, out_dir, writer=(rank % per == 0), decode=not a.latent_only,
                           log=lambda m: print(f"[worker {worker}] {m}", flush=True), sync_export=a.sync_export,
                           export_threads=a.export_threads, metrics_dir=Path(a.measure) if a.measure else None)
        is_w = rank % per == 0
        keys = ("steps", "runs", "denoise_seconds", "decode_seconds", "measure_seconds", "export_encode_seconds",
                "export_stall_seconds")
        t = torch.tensor([stats["seconds"]] + [float(stats[k]) if is_w else 0.0 for k in keys], device="cuda",
                         dtype=torch.float64)
        if world > 1:
            tmax = t.clone()
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            t[0] = tmax[0]
        # the collectives only order the STREAMS: wait for them on the host, so that rank 0 lists the markers after every
        # rank has finished exporting (the first 8-GPU run's manifest missed the last run of seven workers)
        t = t.cpu()
        if rank == 0:
            generate_manifest(out_dir)
            secs = float(t[0])
            tot = dict(zip(keys, [float(v) for v in t[1:]]))
            print(json.dumps({"sweep_runs": int(tot["runs"]), "denoise_steps": int(tot["steps"]), "seconds": secs,
                              "steps_per_s": tot["steps"] / secs if secs else None,
                              "videos_per_hour": tot["runs"] * 3600 / secs if secs else None, "n_gpus": world,
                              "workers": n_workers, "gpus_per_worker": per,
                              "layout": "pairs" if per == 2 else "replicas", "decode": not a.latent_only,
                              "export": "synchronous" if a.sync_export else f"overlapped ({a.export_threads} threads)",
                              "host_cores": os.cpu_count(),
                              "sum_over_workers": {k: round(v, 3) for k, v in tot.items() if k.endswith("seconds")},
                              "what": "wall clock of the slowest worker from its first video to its last marker on disk: "
                                      "denoise loop + VAE decode + uint8 + PNG x16 + GIF + config.json per run"}),
                  flush=True)
        ok = True
    except BaseException:
        import traceback
        traceback.print_exc()
        if world == 1:
            raise
    if world > 1:
        from .dist_util import teardown_process_group
        return teardown_process_group(pipe, ok)
    return 0 if ok else 1


if __name__ == "__main__":
    raise SystemExit(main())
