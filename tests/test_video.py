"""Video path on CPU: deterministic MP4 writer, eager UNet3D pipeline."""
import io

import numpy as np
import pytest
import torch

from arbius_b200.codecs.mp4 import demux_samples, encode_jpeg, encode_mp4, parse_boxes


def _frames(n=6, h=48, w=80):
    yy, xx = np.mgrid[0:h, 0:w]
    base = np.stack([(xx * 3) % 256, (yy * 5) % 256, (xx + yy) % 256], -1).astype(np.uint8)
    return np.stack([np.roll(base, 3 * i, axis=1) for i in range(n)])


def test_jpeg_encoder_decodes_with_pil_and_is_deterministic():
    import PIL.Image
    f = _frames()[0]
    j = encode_jpeg(f, 90)
    assert j[:2] == b"\xff\xd8" and j[-2:] == b"\xff\xd9" and j == encode_jpeg(f, 90)
    dec = np.array(PIL.Image.open(io.BytesIO(j)).convert("RGB"))
    assert dec.shape == f.shape and np.abs(dec.astype(int) - f.astype(int)).mean() < 3.0
    odd = _frames(1, 37, 51)[0]                      # sizes that are not multiples of 8
    dec = np.array(PIL.Image.open(io.BytesIO(encode_jpeg(odd))).convert("RGB"))
    assert dec.shape == odd.shape


def test_mp4_structure_and_round_trip():
    fr = _frames()
    m = encode_mp4(fr, 8)
    assert m == encode_mp4(fr, 8)
    tags = [t for t, _, _ in parse_boxes(m)]
    assert tags == [b"ftyp", b"mdat", b"moov"]
    samples = demux_samples(m)
    assert len(samples) == len(fr) and samples[2] == encode_jpeg(fr[2])
    assert encode_mp4(fr, 24) != m                   # fps is part of the bytes


def test_mp4_plays_in_opencv_if_available(tmp_path):
    cv2 = pytest.importorskip("cv2")
    fr = _frames(5)
    p = tmp_path / "out-1.mp4"
    p.write_bytes(encode_mp4(fr, 8))
    cap = cv2.VideoCapture(str(p))
    n = 0
    while True:
        ok, img = cap.read()
        if not ok:
            break
        n += 1
    if n == 0:
        pytest.skip("this OpenCV build has no MJPEG-in-MP4 demuxer")
    assert n == 5 and abs(cap.get(cv2.CAP_PROP_FPS) - 8) < 1e-3


def test_eager_video_pipeline_cpu():
    from arbius_b200.models.video.pipeline import TextToVideoConfig, TextToVideoPipeline
    pipe = TextToVideoPipeline(TextToVideoConfig.tiny(), backend="eager", device="cpu", dtype=torch.float32)
    a = pipe.generate("a cat walking", "", 4, 2, 32, 32, 9.0, "DDIM", seed=7)
    b = pipe.generate("a cat walking", "", 4, 2, 32, 32, 9.0, "DDIM", seed=7)
    assert a.shape == (4, 32, 32, 3) and a.dtype == np.uint8 and (a == b).all()
    mp4 = pipe.predict({"prompt": "x", "num_frames": 4, "num_inference_steps": 2, "width": 32, "height": 32, "fps": 8, "seed": 1})
    assert mp4[4:8] == b"ftyp" and len(demux_samples(mp4)) == 4
    # the queue API gives the same bytes, in order
    q = [{"prompt": "x", "num_frames": 4, "num_inference_steps": 2, "width": 32, "height": 32, "fps": 8, "seed": sd}
         for sd in (1, 2, 1)]
    many = pipe.predict_many(q)
    assert many[0] == mp4 and many[2] == mp4 and many[1] != mp4


def test_frame_split_rules():
    from arbius_b200.models.video.pipeline import frame_split_ok
    assert all(frame_split_ok(24, 72, 128, p) for p in (2, 4, 8))          # ZeroScope XL default clip
    assert all(frame_split_ok(16, 32, 32, p) for p in (2, 4, 8))           # DAMO default clip
    assert not frame_split_ok(10, 72, 128, 4) and not frame_split_ok(24, 72, 128, 3)
    assert not frame_split_ok(200, 72, 128, 8)                              # more frames than one attention tile holds
    assert frame_split_ok(8, 4, 4, 2) and not frame_split_ok(2, 3, 3, 2)   # 1 frame x 9 px does not split into 4 shards
