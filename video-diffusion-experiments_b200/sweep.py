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
    """frames/frame_%04d.png, <id>.gif, then config.json LAST (it is the completion marker) -- 05:172-190."""
    exp_dir = Path(output_dir) / run.experiment_id
    frames_dir = exp_dir / "frames"
    frames_dir.mkdir(parents=True, exist_ok=True)
    if frames is not None:
        for i, fr in enumerate(frames):
            fr.save(frames_dir / f"frame_{i:04d}.png")
        frames[0].save(os.fspath(exp_dir / f"{run.experiment_id}.gif"), save_all=True, append_images=frames[1:],
                       optimize=False, duration=100, loop=0)
    tmp = exp_dir / "config.json.tmp"
    with open(tmp, "w") as f:
        json.dump(asdict(run), f, indent=2)
    os.replace(tmp, exp_dir / "config.json")
    return exp_dir


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
               log=print) -> dict:
    """Walk one worker's list the way the reference's loops do (generate -> save -> marker).  In a CFG pair
    both ranks call the pipeline (they stay in lock-step through the per-step all-gather); only `writer` saves."""
    import torch
    done_steps, done_runs, t0 = 0, 0, time.perf_counter()
    for run in my_runs:
        if is_done(run, output_dir):
            log(f"  Skipping {run.experiment_id} (already exists)")
            continue
        generator = torch.Generator().manual_seed(run.seed)          # 05:156 torch.manual_seed(seed): CPU generator
        out = pipe(prompt=run.prompt, negative_prompt=run.negative_prompt, num_frames=run.num_frames,
                   guidance_scale=run.guidance_scale, num_inference_steps=run.num_inference_steps,
                   height=run.height, width=run.width, generator=generator,
                   output_type="pil" if decode else "latent")
        if writer:
            save_experiment(out.frames[0] if decode else None, run, output_dir)
        done_steps += run.num_inference_steps
        done_runs += 1
    torch.cuda.synchronize()
    return {"runs": done_runs, "steps": done_steps, "seconds": time.perf_counter() - t0}


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
    stats = run_worker(pipe, mine, out_dir, writer=(rank % per == 0), decode=not a.latent_only,
                       log=lambda m: print(f"[worker {worker}] {m}", flush=True))
    t = torch.tensor([stats["seconds"], float(stats["steps"]) if rank % per == 0 else 0.0,
                      float(stats["runs"]) if rank % per == 0 else 0.0], device="cuda", dtype=torch.float64)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        t[0] = tmax[0]
    if rank == 0:
        generate_manifest(out_dir)
        secs, steps, nruns = float(t[0]), float(t[1]), float(t[2])
        print(json.dumps({"sweep_runs": int(nruns), "denoise_steps": int(steps), "seconds": secs,
                          "steps_per_s": steps / secs if secs else None,
                          "videos_per_hour": nruns * 3600 / secs if secs else None, "n_gpus": world,
                          "workers": n_workers, "gpus_per_worker": per}), flush=True)
    if world > 1:
        # captured graphs hold the pair communicator: drop them and leave without NCCL's finalizers (see bench.py)
        for pl in pipe._plans.values():
            pl.graph = None
        torch.cuda.synchronize()
        dist.barrier()
        import sys
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
