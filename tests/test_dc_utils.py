"""utils.dc_utils keeps the reference's signatures (utils/dc_utils.py:18-86) so run.py imports resolve."""
import inspect
import os

import cv2
import numpy as np


def test_signatures_match_reference():
    from utils import dc_utils

    assert list(inspect.signature(dc_utils.read_video_frames).parameters) == ["video_path", "process_length", "target_fps", "max_res"]
    assert list(inspect.signature(dc_utils.save_video).parameters) == ["frames", "output_video_path", "fps", "is_depths", "grayscale"]


def test_read_video_frames_stride_length_and_max_res(tmp_path):
    from utils.dc_utils import inferno_lut, read_video_frames

    path = os.path.join(tmp_path, "clip.mp4")
    w = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*"mp4v"), 24.0, (64, 48))
    for t in range(12):
        w.write(np.full((48, 64, 3), t * 20, np.uint8))
    w.release()
    frames, fps = read_video_frames(path, -1, -1, -1)
    assert frames.shape == (12, 48, 64, 3) and frames.dtype == np.uint8 and fps == 24.0
    assert abs(int(frames[5].mean()) - 100) <= 8  # mp4v is lossy
    frames, fps = read_video_frames(path, 6, 12, 32)       # first 6 source frames, every 2nd, longest side 32
    assert frames.shape == (3, 24, 32, 3) and fps == 12
    lut = inferno_lut()
    assert lut.shape == (256, 3) and lut.dtype == np.uint8 and lut[0].sum() < 10 and lut[-1][0] > 240


def test_inferno_table_is_pinned():
    """The colour table is OpenCV's COLORMAP_INFERNO, committed as a fixture so a cv2 upgrade cannot change the
    product's bytes silently (matplotlib, the reference's source of the table, is unavailable on this image)."""
    from utils.dc_utils import inferno_lut

    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "inferno_lut_cv2.npy"))
    np.testing.assert_array_equal(inferno_lut(), gold)
