"""Drop-in import path of the reference (`from video_depth_anything.video_depth import
VideoDepthAnything`, run.py:19): thin re-exports of the vda_b200 implementation."""
