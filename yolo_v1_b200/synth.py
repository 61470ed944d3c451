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
