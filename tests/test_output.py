"""The reference's output stage (include/rt_output.h): per-batch 8-bit quantisation, average_images, PPM / PNG files —
checked byte for byte against the reference's own color.h code (oracle/_ref) and against a numpy restatement."""
import struct
import zlib

import numpy as np
import pytest

import raytracing_gpu_b200 as rt
from tests import helpers


def numpy_average(frames):
    """color.h:125-165 restated: (v*v)/(255*255) per byte, float32 sum in frame order, /n, sqrt, clamp, *256."""
    acc = np.zeros(frames.shape[1:], np.float32)
    for f in frames:
        v = f.astype(np.int32)
        acc = acc + (v * v).astype(np.float32) / np.float32(255.0 * 255.0)
    m = np.sqrt(acc / np.float32(len(frames)), dtype=np.float32)
    return (np.float32(256) * np.clip(m, np.float32(0), np.float32(0.999))).astype(np.int32).astype(np.uint8)


def decode_png(path):
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w = 8, b"", None
    while pos < len(data):
        n, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        crc, = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
        assert zlib.crc32(typ + body) == crc
        if typ == b"IHDR":
            w, h, depth, ctype, comp, flt, inter = struct.unpack(">IIBBBBB", body)
            assert (depth, ctype, comp, flt, inter) == (8, 2, 0, 0, 0)
        elif typ == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 3 * w)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 3)


def random_frames(n, h, w, seed):
    rng = np.random.default_rng(seed)
    f = rng.random((n, h, w, 3), dtype=np.float32) ** 2 * 1.3          # some values above 1 (clamped), plenty of dark ones
    f[0, 0, 0] = 0.0
    return f


@pytest.mark.parametrize("n", [1, 2, 5])
def test_average_frames_matches_numpy_restatement(n):
    lin = random_frames(n, 37, 53, seed=n)
    q = np.stack([rt.quantize_frame(f) for f in lin])
    np.testing.assert_array_equal(rt.average_frames(q), numpy_average(q))


@pytest.mark.parametrize("n", [1, 2])
def test_output_pipeline_matches_reference_color_h(n, tmp_path):
    """write_frame_buffer + average_images run from the reference's own sources (oracle/_ref/ref_output_tool); with
    <= 2 frames the float sum does not depend on the directory iteration order the reference uses, so the bytes must
    be identical."""
    h, w = 41, 64
    lin = random_frames(n, h, w, seed=10 + n)
    res = helpers.ref_output_pipeline(lin, str(tmp_path))
    if res is None:
        pytest.skip("oracle/_ref/ref_output_tool not built")
    ref, din = res
    q = np.stack([rt.quantize_frame(f) for f in lin])
    np.testing.assert_array_equal(rt.average_frames(q), ref)
    # the text PPM of one batch is the file the reference wrote, byte for byte
    rt.write_ppm(str(tmp_path / "mine.ppm"), q[0])
    assert open(tmp_path / "mine.ppm", "rb").read() == open(din + "/0.ppm", "rb").read()


def test_output_pipeline_many_frames_close_to_reference(tmp_path):
    n, h, w = 6, 32, 48
    lin = random_frames(n, h, w, seed=3)
    res = helpers.ref_output_pipeline(lin, str(tmp_path))
    if res is None:
        pytest.skip("oracle/_ref/ref_output_tool not built")
    ref, _ = res
    mine = rt.average_frames(np.stack([rt.quantize_frame(f) for f in lin]))
    diff = np.abs(mine.astype(int) - ref.astype(int))
    assert diff.max() <= 1 and (diff > 0).mean() < 1e-3      # only the float summation order may differ


def test_png_round_trip(tmp_path):
    rng = np.random.default_rng(5)
    for h, w in [(1, 1), (7, 5), (300, 211)]:              # the last one needs several stored deflate blocks
        px = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        rt.write_png(str(tmp_path / "a.png"), px)
        np.testing.assert_array_equal(decode_png(tmp_path / "a.png"), px)


def test_bad_arguments():
    with pytest.raises(rt.RtError):
        rt.write_png("/nonexistent_dir/x.png", np.zeros((2, 2, 3), np.uint8))
    sc = rt.Scene()
    with pytest.raises(rt.RtError):
        sc.draw(0, 0.0, 1, 1, image_height=0)               # neither a height nor an aspect ratio


def test_write_npy_is_read_back_by_numpy(tmp_path):
    rng = np.random.default_rng(8)
    for w, h in ((1, 1), (7, 5), (33, 2)):
        img = rng.random((h, w, 3), dtype=np.float32) * 3 - 1
        path = str(tmp_path / f"f_{w}x{h}.npy")
        rt.write_npy(path, img)
        back = np.load(path)
        assert back.dtype == np.float32 and back.shape == (h, w, 3) and (back == img).all()
    with pytest.raises(rt.RtError, match="cannot open"):
        rt.write_npy(str(tmp_path / "no_such_dir" / "x.npy"), img)
