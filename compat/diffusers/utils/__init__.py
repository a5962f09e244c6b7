"""``diffusers.utils`` names the reference imports (05_grid_search_ablation.py:28: export_to_gif)."""
import os
import tempfile


def export_to_gif(image, output_gif_path=None, fps: int = 10):
    if output_gif_path is None:
        output_gif_path = tempfile.NamedTemporaryFile(suffix=".gif").name
    image[0].save(os.fspath(output_gif_path), save_all=True, append_images=image[1:], optimize=False,
                  duration=1000 // fps, loop=0)
    return output_gif_path
