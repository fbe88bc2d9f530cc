"""nvJPEG decode (acmmp_decode_jpeg) against OpenCV's decoder, and the command line on real
.jpg files.  pytest -m gpu."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("no CUDA device", allow_module_level=True)
cv2 = pytest.importorskip("cv2")

from acmmp_b200 import host  # noqa: E402
from acmmp_b200.scene import make_scene  # noqa: E402


def textured(h, w, seed):
    rng = np.random.default_rng(seed)
    base = cv2.resize(rng.integers(0, 256, size=(h // 8 + 1, w // 8 + 1, 3), dtype=np.uint8), (w, h), interpolation=cv2.INTER_CUBIC)
    return np.clip(base.astype(np.int32) + rng.integers(-12, 13, size=(h, w, 3)), 0, 255).astype(np.uint8)


@pytest.mark.parametrize("quality,sub", [(95, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420), (75, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420),
                                         (95, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444)])
def test_decode_matches_opencv(quality, sub):
    img = textured(243, 325, 1)                      # odd sizes: partial MCUs
    ok, enc = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_QUALITY, quality, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, sub])
    assert ok
    grey_ref = cv2.imdecode(enc, cv2.IMREAD_GRAYSCALE)
    bgr_ref = cv2.imdecode(enc, cv2.IMREAD_COLOR)
    grey = host.decode_jpeg(enc.tobytes(), color=False)
    bgr = host.decode_jpeg(enc.tobytes(), color=True)
    assert grey.shape == grey_ref.shape and bgr.shape == bgr_ref.shape
    dg = np.abs(grey.astype(np.int32) - grey_ref.astype(np.int32))
    db = np.abs(bgr.astype(np.int32) - bgr_ref.astype(np.int32))
    # H7: different IDCT / upsampling implementations; luma within 2 grey levels, mean far below 1
    assert dg.max() <= 2 and dg.mean() <= 0.35, (dg.max(), dg.mean())
    if sub == cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444:
        assert db.max() <= 4 and db.mean() <= 0.6, (db.max(), db.mean())
    else:
        # subsampled chroma: libjpeg-turbo interpolates it ("fancy upsampling"), nvJPEG replicates it, so the
        # colours (used for the PLY colours only, never for matching) differ at chroma edges
        assert db.mean() <= 8.0, (db.max(), db.mean())


def test_grey_jpeg_and_errors():
    img = textured(64, 96, 2)[..., 0]
    ok, enc = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_QUALITY, 90])
    grey = host.decode_jpeg(enc.tobytes())
    assert np.abs(grey.astype(np.int32) - cv2.imdecode(enc, cv2.IMREAD_GRAYSCALE).astype(np.int32)).max() <= 2
    with pytest.raises(Exception):
        host.decode_jpeg(b"not a jpeg at all")
