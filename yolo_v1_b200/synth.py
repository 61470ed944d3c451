"""Seeded synthetic inputs for the head hot path.

All tensors are drawn from a CPU ``torch.Generator`` so the oracle, the golden
fixtures and the CUDA kernels see identical bits; callers move them to the GPU.
Target tensors follow the layout the reference's encoder writes
(/root/reference/src/yolo/dataset.py:487-532): slot 0 = (x_cell, y_cell, w, h, 1.0),
slot 1 zeros, one-hot class, all-zero cells elsewhere.
"""
from __future__ import annotations

import torch

PROFILES = ("uniform", "realistic", "ties", "rawhead", "correlated")

# README.md:56-60 of the reference: share of small / medium / large objects on VOC07 test
_SIZE_MIX = (0.066, 0.31, 0.624)
_SMALL_A = (32 / 448) ** 2
_MEDIUM_A = (96 / 448) ** 2


def _gen(seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return g


def make_targets(n: int, seed: int = 1234, S: int = 7, B: int = 2, C: int = 20,
                 p_obj: float = 0.055, size_mix: bool = False, both_slots: float = 0.0) -> torch.Tensor:
    """(n,S,S,B*5+C) fp32 targets; each cell holds an object with probability p_obj
    (0.055 * 49 = 2.7 objects per image, the VOC07-test mean)."""
    g = _gen(seed)
    D = B * 5 + C
    t = torch.zeros(n, S, S, D)
    obj = torch.rand(n, S, S, generator=g) < p_obj
    xy = torch.rand(n, S, S, 2, generator=g)
    if size_mix:
        u = torch.rand(n, S, S, generator=g)
        area_lo = torch.where(u < _SIZE_MIX[0], torch.tensor(0.0004),
                              torch.where(u < _SIZE_MIX[0] + _SIZE_MIX[1], torch.tensor(_SMALL_A),
                                          torch.tensor(_MEDIUM_A)))
        area_hi = torch.where(u < _SIZE_MIX[0], torch.tensor(_SMALL_A),
                              torch.where(u < _SIZE_MIX[0] + _SIZE_MIX[1], torch.tensor(_MEDIUM_A),
                                          torch.tensor(0.81)))
        area = area_lo + (area_hi - area_lo) * torch.rand(n, S, S, generator=g)
        aspect = torch.exp((torch.rand(n, S, S, generator=g) - 0.5) * 1.2)
        w = torch.sqrt(area * aspect).clamp(max=1.0)
        h = torch.sqrt(area / aspect).clamp(max=1.0)
        wh = torch.stack([w, h], dim=-1)
    else:
        wh = 0.02 + 0.88 * torch.rand(n, S, S, 2, generator=g)
    cls = torch.randint(0, C, (n, S, S), generator=g)
    t[..., 0:2] = xy
    t[..., 2:4] = wh
    t[..., 4] = 1.0
    t[..., B * 5:] = torch.nn.functional.one_hot(cls, C).float()
    if both_slots > 0 and B > 1:
        second = torch.rand(n, S, S, generator=g) < both_slots
        t[..., 5:9] = torch.where(second[..., None], torch.rand(n, S, S, 4, generator=g),
                                  torch.zeros(()))
        t[..., 9] = second.float()
    t = t * obj[..., None].float()
    return t.contiguous()


def make_preds(n: int, profile: str = "uniform", seed: int = 1234, S: int = 7, B: int = 2,
               C: int = 20, targets: torch.Tensor | None = None) -> torch.Tensor:
    """(n,S,S,B*5+C) fp32 predictions.

    uniform    U[0,1) everywhere (about 68 of 98 candidates pass conf 0.3: NMS stress)
    realistic  conf = U^4, class = softmax(3*randn), xy ~ U, wh ~ U[0.05,0.8]
    ties       every value quantised to multiples of 1/8 (stable-order / >= edges)
    rawhead    0.5*randn (what an untrained head emits; negative sizes, conf > 1 ...)
    correlated realistic noise, and in object cells of ``targets`` box 0 = target +
               N(0,0.05), conf ~ U[0.3,1], true class prob high w.p. 0.8 (so TPs exist)
    """
    g = _gen(seed + 7919)
    D = B * 5 + C
    if profile == "uniform":
        p = torch.rand(n, S, S, D, generator=g)
    elif profile == "ties":
        p = torch.floor(torch.rand(n, S, S, D, generator=g) * 9) / 8
    elif profile == "rawhead":
        p = 0.5 * torch.randn(n, S, S, D, generator=g)
    elif profile in ("realistic", "correlated"):
        p = torch.empty(n, S, S, D)
        for b in range(B):
            p[..., b * 5:b * 5 + 2] = torch.rand(n, S, S, 2, generator=g)
            p[..., b * 5 + 2:b * 5 + 4] = 0.05 + 0.75 * torch.rand(n, S, S, 2, generator=g)
            p[..., b * 5 + 4] = torch.rand(n, S, S, generator=g) ** 4
        p[..., B * 5:] = torch.softmax(3.0 * torch.randn(n, S, S, C, generator=g), dim=-1)
        if profile == "correlated":
            if targets is None:
                raise ValueError("profile 'correlated' needs targets")
            obj = targets[..., 4] > 0
            noisy = targets[..., 0:4] + 0.05 * torch.randn(n, S, S, 4, generator=g)
            noisy = noisy.clamp(0.0, 1.0)
            conf = 0.3 + 0.7 * torch.rand(n, S, S, generator=g)
            hit = torch.rand(n, S, S, generator=g) < 0.8
            sharp = 0.1 * torch.rand(n, S, S, C, generator=g) + 0.85 * targets[..., B * 5:]
            p[..., 0:4] = torch.where(obj[..., None], noisy, p[..., 0:4])
            p[..., 4] = torch.where(obj, conf, p[..., 4])
            p[..., B * 5:] = torch.where((obj & hit)[..., None], sharp, p[..., B * 5:])
    else:
        raise ValueError(f"unknown profile {profile!r}; choose from {PROFILES}")
    return p.contiguous()


# ---------------------------------------------------------------------------------------------
# Bit-reproducible inputs.  The torch generators above go through exp / pow / softmax, whose
# vectorised and scalar-tail code paths differ in the last bit, so their output depends on the
# ATen thread count and the CPU's vector width.  Fixtures that are REGENERATED on another
# machine (tests/golden/map_big.npz stores only a hash of its 58 MB of inputs) are drawn here:
# a counter-based integer hash (splitmix64) turned into k / 2^24 floats, and nothing but
# IEEE-exact operators (+ - * / sqrt, comparisons) after that.
_SM_GAMMA = 0x9E3779B97F4A7C15
_MASK64 = (1 << 64) - 1


def _mix64(z):
    import numpy as np

    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def _mix64_int(z: int) -> int:
    z &= _MASK64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK64
    return z ^ (z >> 31)


def exact_uniform(shape, seed: int, stream: int):
    """float32 array of k / 2^24, k = top 24 bits of splitmix64(hash(seed, stream) + gamma * index)."""
    import numpy as np

    count = int(np.prod(shape))
    base = np.uint64(_mix64_int(int(seed) * 0xD1342543DE82EF95 + int(stream) * 0xA0761D6478BD642F + 1))
    idx = np.arange(count, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = _mix64(base + (idx + np.uint64(1)) * np.uint64(_SM_GAMMA))
    return ((z >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)).reshape(shape)


def make_exact(n: int, seed: int = 1234, S: int = 7, B: int = 2, C: int = 20, p_obj: float = 0.055):
    """(pred, targets) in the spirit of the `correlated` profile with VOC's size mix (SURVEY.md
    section 8d config 3), identical bits on any machine.  Returns torch float32 tensors."""
    import numpy as np

    f = np.float32
    u = lambda shape, stream: exact_uniform(shape, seed, stream)  # noqa: E731
    D = B * 5 + C
    # ---- targets: slot 0 = (x, y, w, h, 1), one-hot class (dataset.py:516-530)
    obj = u((n, S, S), 1) < f(p_obj)
    mix = u((n, S, S), 2)
    small, medium = mix < f(_SIZE_MIX[0]), mix < f(_SIZE_MIX[0] + _SIZE_MIX[1])
    lo = np.where(small, f(0.0004), np.where(medium, f(_SMALL_A), f(_MEDIUM_A))).astype(f)
    hi = np.where(small, f(_SMALL_A), np.where(medium, f(_MEDIUM_A), f(0.81))).astype(f)
    area = lo + (hi - lo) * u((n, S, S), 3)
    aspect = (f(0.55) + u((n, S, S), 4)) / (f(0.55) + u((n, S, S), 5))       # 0.35 .. 2.8
    w = np.minimum(np.sqrt(area * aspect), f(1.0))
    h = np.minimum(np.sqrt(area / aspect), f(1.0))
    cls = np.minimum((u((n, S, S), 6) * f(C)).astype(np.int64), C - 1)
    t = np.zeros((n, S, S, D), f)
    t[..., 0:2] = u((n, S, S, 2), 7)
    t[..., 2] = w
    t[..., 3] = h
    t[..., 4] = f(1.0)
    t[..., B * 5:] = (cls[..., None] == np.arange(C)).astype(f)
    t *= obj[..., None].astype(f)
    # ---- predictions: low-confidence noise everywhere ...
    p = np.empty((n, S, S, D), f)
    for b in range(B):
        p[..., b * 5:b * 5 + 2] = u((n, S, S, 2), 10 + 3 * b)
        p[..., b * 5 + 2:b * 5 + 4] = f(0.05) + f(0.75) * u((n, S, S, 2), 11 + 3 * b)
        c2 = u((n, S, S), 12 + 3 * b)
        c2 = c2 * c2
        p[..., b * 5 + 4] = c2 * c2
    q = u((n, S, S, C), 20)
    q = q * q
    p[..., B * 5:] = q * q * f(0.5)
    # ... and in object cells box 0 = target + noise, confidence 0.3..1, the true class high w.p. 0.8
    noise = (u((n, S, S, 4), 21) + u((n, S, S, 4), 22) + u((n, S, S, 4), 23) + u((n, S, S, 4), 24) - f(2.0)) * f(0.0866)
    noisy = np.clip(t[..., 0:4] + noise, f(0.0), f(1.0))
    conf = f(0.3) + f(0.7) * u((n, S, S), 25)
    hit = obj & (u((n, S, S), 26) < f(0.8))
    sharp = f(0.1) * u((n, S, S, C), 27) + f(0.85) * t[..., B * 5:]
    p[..., 0:4] = np.where(obj[..., None], noisy, p[..., 0:4])
    p[..., 4] = np.where(obj, conf, p[..., 4])
    p[..., B * 5:] = np.where(hit[..., None], sharp, p[..., B * 5:])
    return torch.from_numpy(np.ascontiguousarray(p)), torch.from_numpy(np.ascontiguousarray(t))


def head_linear_inputs(M: int, seed: int):
    """x [M, 4096], w [1470, 4096], b [1470], targets [M, 7, 7, 30] as fp32 numpy arrays for the fused head Linear + loss
    (tests/golden/make_head_linear_golden.py): x and w have 8 significant bits, i.e. are exact in bf16 AND fp16."""
    import numpy as np

    def q(a):
        return np.round(a * 128.0) / 128.0          # multiples of 2^-7 below 1

    x = q(exact_uniform((M, 4096), seed, 11) - 0.5).astype(np.float32)
    w = (q(exact_uniform((1470, 4096), seed, 12) - 0.5) * 2.0 ** -5).astype(np.float32)
    b = (0.15 + 0.5 * exact_uniform((1470,), seed, 13)).astype(np.float32)
    _, tgt = make_exact(M, seed=seed, p_obj=0.12)
    return x, w, b, tgt.numpy()
