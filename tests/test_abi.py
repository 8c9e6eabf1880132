"""CPU-side checks of the C ABI and host logic: the in-tree library builds/loads without a GPU and
exports every function include/vda_b200.h declares; no compute call is made here."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "vda_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vda_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import vda_b200.lib as L

    lib = L.load()
    names = _declared()
    assert len(names) >= 19
    for n in names:
        assert isinstance(getattr(lib, n), ctypes._CFuncPtr), n
    # the ctypes table in lib.py covers the header exactly
    assert sorted(L.exported_symbols()) == names


def test_no_cpu_fallback():
    import torch

    import vda_b200.lib as L

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(L.VdaError):
        L.init()
    from video_depth_anything.video_depth import VideoDepthAnything

    m = VideoDepthAnything(encoder="vits", features=64, out_channels=[48, 96, 192, 384])
    with pytest.raises(L.VdaError):
        m(torch.zeros(1, 2, 3, 28, 28))


def test_argument_errors_are_reported_without_a_gpu():
    import vda_b200.lib as L

    lib = L.load()
    rc = lib.vda_lsq_solve(None, None, 0, None)
    assert rc != 0 and b"bad arguments" in lib.vda_last_error()


def test_state_dict_schema_matches_reference_keys():
    from oracle import refload
    from video_depth_anything.video_depth import VideoDepthAnything
    from video_depth_anything.video_depth_stream import VideoDepthAnything as Stream

    for arch, f, oc in [("vits", 64, [48, 96, 192, 384]), ("vitl", 256, [256, 512, 1024, 1024])]:
        want = refload.state_dict_shapes(arch)
        for cls in (VideoDepthAnything, Stream):
            got = {k: tuple(v.shape) for k, v in cls(encoder=arch, features=f, out_channels=oc).state_dict().items()}
            assert got == want


def test_window_indices_and_input_size_match_oracle():
    from oracle import vda_oracle as O
    from vda_b200.preprocess import net_input_size
    from vda_b200.video_depth import window_frame_indices

    for n in (1, 5, 22, 23, 32, 33, 100, 1024):
        assert window_frame_indices(n) == O.window_frame_indices(n)
    assert len(window_frame_indices(1024)) == 47 and len(window_frame_indices(512)) == 24
    for hw in [(518, 518), (720, 1280), (1080, 1920), (480, 640), (44, 60), (2160, 3840), (100, 400)]:
        assert net_input_size(*hw) == O.net_input_size(*hw)


def test_preprocess_matches_oracle_bitwise():
    from oracle import make_golden as G
    from oracle import vda_oracle as O
    from vda_b200.preprocess import _one

    fr = G.video_frames(2, 44, 60)
    for f in fr:
        dst = np.empty((3, 56, 70), np.float32)
        _one(f, 56, 70, dst)
        np.testing.assert_array_equal(dst, O.preprocess_frame(f, 56, 70))


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="reference tree only exists in the build container")
@pytest.mark.parametrize("script", ["run.py", "run_streaming.py"])
def test_reference_entry_points_import_against_the_repo(script):
    """The reference's own run.py / run_streaming.py, unmodified, resolve every import (`video_depth_anything.*`,
    `utils.dc_utils`) against this repo when it is first on PYTHONPATH: `--help` runs their module level code up to
    argparse (run.py:15-40, run_streaming.py:15-38)."""
    import subprocess
    import sys

    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, os.path.join("/root/reference", script), "--help"], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "--encoder" in r.stdout and "--input_size" in r.stdout
