"""Synthetic stereo pairs for the configs BASELINE.json names.

Three match-distance distributions, because the cost of the search is data dependent:

  * "nat"   value-noise texture; the additional image is the main image warped by a smooth horizontal disparity
            field plus +-4 noise.  README-like: most queries match in ring 0, a long tail reaches far rings.
  * "geo"   main = random colours of a 64-colour lattice palette; an additional pixel is a palette colour with
            probability rho, else a filler colour that matches no palette colour at precision <= 36.
            First-match probe index is geometric: mean ring ~ max_rings/4 with rho = 1/64 at 128 rings.
  * "worst" as "geo" with rho = 0 and one planted pixel (so that max != 0): every query scans every ring.

The generators are written with torch so that the benchmark can build 4K batches on the GPU; on the CPU they
are deterministic for a given seed (tests compare oracle and GPU on the SAME bytes, so cross-device identity of
the generator is not needed).
"""
from __future__ import annotations

import torch

PALETTE_LEVELS = (16, 88, 160, 232)   # spacing 72: two different palette colours never match at precision <= 72
FILLER_LEVELS = (52, 124, 196)        # 36 away from every palette level: never within precision <= 36


def _gen(seed: int, device) -> torch.Generator:
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    return g


def _value_noise(h: int, w: int, cell: int, g: torch.Generator, device) -> torch.Tensor:
    """bilinear value noise in [0,1): random lattice with `cell`-pixel spacing"""
    gh, gw = h // cell + 2, w // cell + 2
    lat = torch.rand((gh, gw), generator=g, device=device, dtype=torch.float32)
    ys = (torch.arange(h, device=device, dtype=torch.float32) + 0.5) / cell
    xs = (torch.arange(w, device=device, dtype=torch.float32) + 0.5) / cell
    y0, x0 = ys.floor().long(), xs.floor().long()
    fy, fx = (ys - y0).unsqueeze(1), (xs - x0).unsqueeze(0)
    fy, fx = fy * fy * (3 - 2 * fy), fx * fx * (3 - 2 * fx)  # smoothstep
    a = lat[y0][:, x0]
    b = lat[y0][:, x0 + 1]
    c = lat[y0 + 1][:, x0]
    d = lat[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def _fractal(h, w, cells, amps, g, device):
    acc = torch.zeros((h, w), device=device, dtype=torch.float32)
    for cell, amp in zip(cells, amps):
        acc += amp * (_value_noise(h, w, cell, g, device) - 0.5)
    return acc


def nat_pair(w: int, h: int, seed: int = 1, device="cpu", max_disp: int = 64):
    """-> (main_rgba [h,w,4] u8, add_rgb [h,w,3] u8)"""
    g = _gen(seed, device)
    lum = _fractal(h, w, (192, 48, 12, 3), (1.0, 0.55, 0.28, 0.10), g, device)
    chans = []
    for _ in range(3):
        chroma = _fractal(h, w, (256, 64), (0.9, 0.35), g, device)
        chans.append(128.0 + 85.0 * lum + 55.0 * chroma)  # channel std ~25, like the README photos
    main = torch.stack(chans, dim=-1)
    noise = torch.randint(-4, 5, (h, w, 3), generator=g, device=device).float()
    main_u8 = (main + noise).clamp(0, 255).to(torch.uint8)
    # additional = main sampled at x + disparity(x, y), smooth field in [0, max_disp]
    disp = (_value_noise(h, w, 160, g, device) * max_disp).round().long()
    xs = torch.arange(w, device=device).unsqueeze(0).expand(h, w)
    src_x = (xs + disp).clamp(0, w - 1)
    add = torch.gather(main, 1, src_x.unsqueeze(-1).expand(h, w, 3))
    noise2 = torch.randint(-4, 5, (h, w, 3), generator=g, device=device).float()
    add_u8 = (add * 0.97 + 2.0 + noise2).clamp(0, 255).to(torch.uint8)
    alpha = torch.full((h, w, 1), 255, dtype=torch.uint8, device=device)
    return torch.cat([main_u8, alpha], dim=-1).contiguous(), add_u8.contiguous()


def _palette(device):
    lv = torch.tensor(PALETTE_LEVELS, dtype=torch.uint8, device=device)
    idx = torch.arange(64, device=device)
    return torch.stack([lv[idx // 16], lv[(idx // 4) % 4], lv[idx % 4]], dim=-1)  # [64,3]


def geo_pair(w: int, h: int, seed: int = 1, device="cpu", rho: float = 1.0 / 64.0, plant: bool = False):
    """-> (main_rgba, add_rgb); P(an additional pixel matches a given query) = rho / 64"""
    g = _gen(seed, device)
    pal = _palette(device)
    main_idx = torch.randint(0, 64, (h, w), generator=g, device=device)
    main = pal[main_idx]
    fl = torch.tensor(FILLER_LEVELS, dtype=torch.uint8, device=device)
    add = fl[torch.randint(0, 3, (h, w, 3), generator=g, device=device)]
    if rho > 0:
        hit = torch.rand((h, w), generator=g, device=device) < rho
        add_idx = torch.randint(0, 64, (h, w), generator=g, device=device)
        add = torch.where(hit.unsqueeze(-1), pal[add_idx], add)
    if plant:
        add[h // 4, w // 4] = pal[0]
    alpha = torch.full((h, w, 1), 255, dtype=torch.uint8, device=device)
    return torch.cat([main, alpha], dim=-1).contiguous(), add.contiguous()


def worst_pair(w: int, h: int, seed: int = 1, device="cpu"):
    return geo_pair(w, h, seed, device, rho=0.0, plant=True)


GENERATORS = {"nat": nat_pair, "geo": geo_pair, "worst": worst_pair}


def make_pair(kind: str, w: int, h: int, seed: int = 1, device="cpu"):
    return GENERATORS[kind](w, h, seed, device)
