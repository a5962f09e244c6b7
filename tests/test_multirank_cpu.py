"""N > 1 host logic on CPU (gloo, world size 2): the CFG-split denoise loop run on the oracle
must equal the single-process loop bit for bit, and the sweep schedule must be the reference's 78 runs."""
import json
import os
import sys
import tempfile

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ad  # noqa: E402,F401
from b200ad import sweep  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden", "sweep_manifest.json")
SMALL = dict(block_out_channels=(32, 64, 64, 64), attention_head_dim=4, motion_num_attention_heads=4,
             norm_num_groups=8)


def _small_cfg():
    from oracle.unet_motion import UNetMotionConfig
    import dataclasses
    names = {f.name for f in dataclasses.fields(UNetMotionConfig)}
    return UNetMotionConfig(**{k: v for k, v in SMALL.items() if k in names})


def _inputs():
    lat = torch.randn(1, 4, 4, 8, 8, generator=torch.Generator().manual_seed(42))
    txt = torch.randn(2, 77, 768, generator=torch.Generator().manual_seed(43))
    return lat, txt


def _split_rank(rank, world, port, out_path, steps, g):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    from oracle.unet_motion import make_oracle_unet
    from oracle.schedulers import OracleDDIM
    unet = make_oracle_unet(seed=0, cfg=_small_cfg())
    lat, txt = _inputs()
    sch = OracleDDIM()
    sch.set_timesteps(steps)
    my_text = txt[rank][None]                       # rank 0 = uncond row, rank 1 = cond row
    with torch.no_grad():
        for t in sch.timesteps:
            eps = unet(lat, t, encoder_hidden_states=my_text).sample.contiguous()
            both = [torch.empty_like(eps) for _ in range(world)]
            dist.all_gather(both, eps)              # the one exchange per denoise step
            guided = both[0] + g * (both[1] - both[0])
            lat = sch.step(guided, t, lat)          # redundantly on both ranks: latents stay identical
    torch.save(lat, f"{out_path}.{rank}")
    dist.barrier()
    dist.destroy_process_group()


def test_cfg_split_loop_equals_single_process_loop_bitwise():
    from oracle.unet_motion import make_oracle_unet
    from oracle.schedulers import OracleDDIM, denoise_loop
    steps, g = 3, 9.0
    torch.set_num_threads(2)
    lat, txt = _inputs()
    # single process; each CFG half evaluated as its own batch-1 call so that CPU GEMM blocking is identical
    unet = make_oracle_unet(seed=0, cfg=_small_cfg())
    sch = OracleDDIM()
    sch.set_timesteps(steps)
    ref = lat
    with torch.no_grad():
        for t in sch.timesteps:
            u = unet(ref, t, encoder_hidden_states=txt[0:1]).sample
            c = unet(ref, t, encoder_hidden_states=txt[1:2]).sample
            ref = sch.step(u + g * (c - u), t, ref)
    # the batched (B=2) loop of the pipeline agrees to rounding: the halves are independent through the UNet
    batched = denoise_loop(make_oracle_unet(seed=0, cfg=_small_cfg()), OracleDDIM(), lat, txt, g, steps)
    assert ((batched - ref).norm() / ref.norm()).item() < 1e-5
    with tempfile.TemporaryDirectory() as d:
        out = os.path.join(d, "lat")
        port = 29600 + os.getpid() % 300
        mp.spawn(_split_rank, args=(2, port, out, steps, g), nprocs=2, join=True)
        l0, l1 = torch.load(out + ".0"), torch.load(out + ".1")
    assert torch.equal(l0, l1), "both ranks of a CFG pair must hold bit-identical latents"
    assert torch.equal(l0, ref), "split loop must equal the single-process loop bit for bit"


def test_sweep_is_the_references_78_runs():
    gold = json.load(open(GOLDEN))
    runs = sweep.sweep_runs()
    assert len(runs) == 78 and sum(r.num_inference_steps for r in runs) == 2130
    mine = sorted([r.experiment_id, r.video_name, r.guidance_scale, r.num_inference_steps, r.phase] for r in runs)
    assert mine == sorted(gold["experiments"])
    assert gold["grid_params"] == {"cfg_values": sweep.CFG_VALUES, "steps_values": sweep.STEPS_VALUES,
                                   "default_cfg": sweep.DEFAULT_CFG, "default_steps": sweep.DEFAULT_STEPS}
    assert all(r.seed == 42 and r.num_frames == 16 and r.height == r.width == 512 for r in runs)


@pytest.mark.parametrize("n_workers,bound", [(1, 2130), (2, 1065), (4, 540), (8, 270)])
def test_lpt_assignment_is_a_balanced_partition(n_workers, bound):
    runs = sweep.sweep_runs()
    parts = sweep.assign_lpt(runs, n_workers)
    ids = [r.experiment_id for p in parts for r in p]
    assert sorted(ids) == sorted(r.experiment_id for r in runs) and len(set(ids)) == 78
    loads = [sum(r.num_inference_steps for r in p) for p in parts]
    assert max(loads) <= bound, loads                      # ideal 2130 / n_workers (532.5 per pair at 8 GPUs); LPT is within one 15-step run of it
    assert parts == sweep.assign_lpt(runs, n_workers)      # deterministic: every rank computes the same lists


def test_resume_markers_and_manifest(tmp_path):
    runs = sweep.sweep_runs(video_filter="corgi")
    assert len(runs) == 13
    for r in runs[:5]:
        sweep.save_experiment(None, r, tmp_path)
    assert [sweep.is_done(r, tmp_path) for r in runs] == [True] * 5 + [False] * 8
    cfg = json.load(open(tmp_path / runs[0].experiment_id / "config.json"))
    assert list(cfg) == ["experiment_id", "video_name", "prompt", "negative_prompt", "guidance_scale",
                         "num_inference_steps", "phase", "seed", "num_frames", "height", "width"]
    man = sweep.generate_manifest(tmp_path)
    assert len(man["experiments"]) == 5 and set(man["experiments"][0]) == {"experiment_id", "video_name", "cfg",
                                                                          "steps", "phase"}
    assert sweep.worker_layout(8, True) == (4, 2) and sweep.worker_layout(8, False) == (8, 1)
    assert sweep.worker_layout(1, True) == (1, 1)
