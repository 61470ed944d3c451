"""Image preprocessing: oracle/preprocess_oracle.py and yh_preprocess_rgb8 against the reference's
transform (src/yolo/inference.py:57-66: torchvision Resize((448, 448)) on a PIL image, ToTensor, Normalize).
Golden hashes come from the real torchvision + Pillow (tests/golden/make_preprocess_golden.py); where the two
libraries are importable (they are part of this image) the comparison is also made live.  Bit-exact throughout."""
import hashlib
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
from make_preprocess_golden import CASES, make_image  # noqa: E402

from oracle import preprocess_oracle as po  # noqa: E402

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "preprocess.npz"))
SMALL = [i for i, (h, w) in enumerate(CASES) if h * w <= 400_000]   # the pure-numpy oracle is slow on the largest


def _sha(a: np.ndarray) -> bytes:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest()


@pytest.mark.parametrize("i", range(len(CASES)))
def test_oracle_matches_golden(i):
    h, w = CASES[i]
    img = make_image(100 + i, h, w)
    assert _sha(img) == GOLD[f"in_sha/{i}"].tobytes(), "seeded image differs from the one the golden was made from"
    got = po.preprocess(img)
    assert got.dtype == np.float32 and got.shape == (3, 448, 448)
    assert np.array_equal(got[:, ::37, ::41].view(np.uint32), GOLD[f"sample/{i}"].view(np.uint32))
    assert _sha(got) == GOLD[f"out_sha/{i}"].tobytes()


def test_oracle_matches_live_torchvision():
    transforms = pytest.importorskip("torchvision.transforms")
    Image = pytest.importorskip("PIL.Image")
    tf = transforms.Compose([transforms.Resize((448, 448)), transforms.ToTensor(),
                             transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    rng = np.random.default_rng(7)
    for _ in range(6):
        h, w = int(rng.integers(1, 700)), int(rng.integers(1, 700))
        img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
        want = tf(Image.fromarray(img)).numpy()
        assert np.array_equal(po.preprocess(img).view(np.uint32), want.view(np.uint32)), (h, w)


def test_coefficients_known_answers():
    # same size: identity taps (1, 0); 2x downscale: taps 1/8 3/8 3/8 1/8 at full weight inside the image
    b, k = po.precompute_coeffs(448, 448)
    assert (b[:, 0] == np.arange(448)).all() and (k[:, 0] == 1 << 22).all() and (k[:, 1:] == 0).all()
    b, k = po.precompute_coeffs(896, 448)
    assert tuple(b[5]) == (9, 4) and list(k[5][:4]) == [524288, 1572864, 1572864, 524288]
    assert int(k[5].sum()) == 1 << 22


# ------------------------------------------------------------------------------------------ GPU
gpu = pytest.mark.gpu


@gpu
@pytest.mark.parametrize("i", range(len(CASES)))
def test_cuda_matches_golden_and_oracle(i):
    from yolo_v1_b200 import preprocess
    h, w = CASES[i]
    img = make_image(100 + i, h, w)
    got = preprocess.preprocess_rgb8(torch.from_numpy(img).cuda()).cpu().numpy()
    assert _sha(got) == GOLD[f"out_sha/{i}"].tobytes()
    if i in SMALL:
        assert np.array_equal(got.view(np.uint32), po.preprocess(img).view(np.uint32))


@gpu
def test_cuda_matches_live_torchvision_random_sizes_and_other_targets():
    transforms = pytest.importorskip("torchvision.transforms")
    Image = pytest.importorskip("PIL.Image")
    from yolo_v1_b200 import preprocess
    rng = np.random.default_rng(11)
    for r in range(42):
        h, w = int(rng.integers(1, 1500)), int(rng.integers(1, 1500))
        if r >= 40:   # a span of input pixels too long for the staged horizontal pass: the unstaged kernel
            h, w = 3, 61_000 + r
        oh, ow = ((448, 448), (224, 224), (300, 517))[r % 3]
        tf = transforms.Compose([transforms.Resize((oh, ow)), transforms.ToTensor(),
                                 transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
        img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
        want = tf(Image.fromarray(img)).numpy()
        got = preprocess.preprocess_rgb8(torch.from_numpy(img).cuda(), (oh, ow)).cpu().numpy()
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (h, w, oh, ow)


@gpu
def test_engine_preprocess_image_equals_its_transform():
    Image = pytest.importorskip("PIL.Image")
    from yolo_v1_b200 import YOLOInference

    class Head(torch.nn.Module):
        S, B = 7, 2

        def forward(self, x):
            return torch.zeros(x.shape[0], 7, 7, 30, device=x.device)

    eng = YOLOInference(Head(), device="cuda")
    img = Image.fromarray(make_image(5, 333, 500))
    got = eng.preprocess_image(img)
    want = eng.transform(img).unsqueeze(0)
    assert got.is_cuda and got.shape == (1, 3, 448, 448) and torch.equal(got.cpu(), want)
    gray = Image.fromarray(make_image(6, 64, 48)[..., 0])           # a mode-L image is converted like load_image does
    assert torch.equal(eng.preprocess_image(gray).cpu(), eng.transform(gray.convert("RGB")).unsqueeze(0))


@gpu
def test_bad_arguments():
    from yolo_v1_b200 import _lib, preprocess
    with pytest.raises(_lib.YoloHeadError):
        preprocess.preprocess_rgb8(torch.zeros(4, 4, 3, dtype=torch.uint8))          # host tensor: no CPU path
    with pytest.raises(TypeError):
        preprocess.preprocess_rgb8(torch.zeros(4, 4, 3, device="cuda"))               # not uint8


@gpu
def test_predict_images_equals_predict_per_image(tmp_path):
    Image = pytest.importorskip("PIL.Image")
    from yolo_v1_b200 import YOLOInference

    class Head(torch.nn.Module):          # a stand-in backbone: predictions are a fixed function of the input tensor
        S, B = 7, 2

        def forward(self, x):
            feat = torch.nn.functional.adaptive_avg_pool2d(x, (7, 7)).permute(0, 2, 3, 1)     # (N, 7, 7, 3)
            return torch.sigmoid(torch.cat([feat * (i + 1) for i in range(10)], dim=-1))      # (N, 7, 7, 30) in (0, 1)

    eng = YOLOInference(Head(), device="cuda")
    paths = []
    for i, (h, w) in enumerate([(120, 160), (333, 500), (448, 448), (64, 31), (500, 375)]):
        p = tmp_path / f"img{i}.png"
        Image.fromarray(make_image(40 + i, h, w)).save(p)
        paths.append(str(p))
    many = eng.predict_images(paths, conf_threshold=0.3, batch_size=2)
    assert len(many) == len(paths) and sum(len(d) for d in many) > 0
    for p, dets in zip(paths, many):
        assert dets == eng.predict(p, conf_threshold=0.3)


@gpu
@pytest.mark.parametrize("shape", [(5, 375, 500), (3, 448, 448), (4, 37, 1201), (2, 900, 448)])
def test_batch_of_equal_sized_images_equals_per_image(shape):
    from yolo_v1_b200 import preprocess
    n, h, w = shape
    imgs = np.stack([make_image(60 + i, h, w) for i in range(n)])
    d = torch.from_numpy(imgs).cuda()
    batch = preprocess.preprocess_rgb8(d)
    assert batch.shape == (n, 3, 448, 448)
    for i in range(n):
        assert torch.equal(batch[i], preprocess.preprocess_rgb8(d[i]))
    assert np.array_equal(batch[n - 1].cpu().numpy().view(np.uint32), po.preprocess(imgs[n - 1]).view(np.uint32))
