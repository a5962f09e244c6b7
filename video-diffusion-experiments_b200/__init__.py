"""B200-native AnimateDiff denoise step (importable as ``b200ad``; see b200ad.py at the repo root).

The directory name follows the build contract (video-diffusion-experiments_b200/); since it is not a valid Python
identifier the repo-root module ``b200ad.py`` registers this package under the name ``b200ad``.
"""
from . import lib  # noqa: F401

__all__ = ["lib"]
__version__ = "0.1.0"
