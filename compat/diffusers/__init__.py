"""Import-compatible ``diffusers`` shim: put ``<repo>/compat`` on PYTHONPATH and the reference's scripts
(``from diffusers import AnimateDiffPipeline, MotionAdapter, DDIMScheduler, EulerDiscreteScheduler``,
/root/reference/experiments/05_grid_search_ablation.py:27-28) run unchanged on the b200ad path.
Real diffusers is not installable on this image; only the names the reference imports are provided."""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
import b200ad  # noqa: E402,F401
from b200ad.pipeline import AnimateDiffPipeline, AnimateDiffPipelineOutput, MotionAdapter  # noqa: E402,F401
from b200ad.schedulers import DDIMScheduler, EulerDiscreteScheduler  # noqa: E402,F401
from b200ad.unet import UNetMotionModel  # noqa: E402,F401
from b200ad.attention_processor import B200AttnProcessor as AttnProcessor2_0  # noqa: E402,F401

__version__ = "0.31.0+b200ad"
