"""CPU oracle for the AnimateDiff denoise step (TEST INFRASTRUCTURE, not product code).

A plain-PyTorch fp32 restatement of the diffusers modules the reference drives at
/root/reference/experiments/05_grid_search_ablation.py:158-167 (AnimateDiffPipeline ->
UNetMotionModel -> scheduler.step).  diffusers itself (reference requirements.txt:6,
"diffusers>=0.25.0", unpinned, un-vendored) is not installable in this image, so the
restatement follows SURVEY.md Appendix A/B and is validated against the reference's own
structural known-answers (docs/02_video_diffusion_architecture.md:86-91).

PARITY UNPINNED: the reference holds no test, golden vector or stored tensor for this path
(SURVEY.md section 8c); only structural counts and the scheduler formulas are pinned.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  The product (video-diffusion-experiments_b200/) never does.
"""
