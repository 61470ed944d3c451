"""Golden vectors for the image preprocessing (yh_preprocess_rgb8 / oracle/preprocess_oracle.py), produced by the
real thing: the reference's transform (src/yolo/inference.py:57-66) built from the torchvision and Pillow installed
in this container, applied to seeded random RGB images.  Stored per case: the seed and shape (the image is
regenerated from them; its sha256 guards against generator drift), the sha256 of the float32 output bytes and a
strided sample of the output.

    python tests/golden/make_preprocess_golden.py
"""
import hashlib
import os

import numpy as np
from PIL import Image
from torchvision import transforms

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = [(375, 500), (500, 333), (448, 448), (80, 100), (768, 1024), (1, 1), (7, 3000), (448, 200), (300, 448),
         (1200, 1600), (2, 2), (449, 447), (3, 5), (2048, 31)]


def make_image(seed: int, h: int, w: int) -> np.ndarray:
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
    if seed % 3 == 0:   # smooth gradients as well as noise: rounding ties of the fixed-point taps behave differently
        yy, xx = np.mgrid[0:h, 0:w]
        img[..., 0] = (xx * 255 // max(w - 1, 1)).astype(np.uint8)
        img[..., 1] = (yy * 255 // max(h - 1, 1)).astype(np.uint8)
    return img


def main():
    tf = transforms.Compose([transforms.Resize((448, 448)), transforms.ToTensor(),
                             transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    out = {"shapes": np.array(CASES, np.int32)}
    for i, (h, w) in enumerate(CASES):
        img = make_image(100 + i, h, w)
        t = tf(Image.fromarray(img)).numpy()
        assert t.dtype == np.float32 and t.shape == (3, 448, 448)
        out[f"in_sha/{i}"] = np.frombuffer(hashlib.sha256(img.tobytes()).digest(), np.uint8)
        out[f"out_sha/{i}"] = np.frombuffer(hashlib.sha256(t.tobytes()).digest(), np.uint8)
        out[f"sample/{i}"] = t[:, ::37, ::41].copy()
    np.savez_compressed(os.path.join(HERE, "preprocess.npz"), **out)
    print("wrote preprocess.npz", os.path.getsize(os.path.join(HERE, "preprocess.npz")), "bytes")


if __name__ == "__main__":
    main()
