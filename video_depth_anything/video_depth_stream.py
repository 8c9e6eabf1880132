"""Same import path as the reference's video_depth_anything/video_depth_stream.py
(run_streaming.py:21); the implementation lives in vda_b200.video_depth_stream."""
from vda_b200.video_depth_stream import VideoDepthAnything  # noqa: F401
