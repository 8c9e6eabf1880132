"""Same import path as the reference's video_depth_anything/video_depth.py (run.py:19, app.py:20,
benchmark/infer/infer.py:9); the implementation lives in vda_b200.video_depth."""
from vda_b200.video_depth import (INFER_LEN, INTERP_LEN, KEYFRAMES, OVERLAP,  # noqa: F401
                                  VideoDepthAnything)
