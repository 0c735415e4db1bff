"""Synthetic clouds of BASELINE.json's configs, generated with torch so the same code
runs on the GPU (bench inputs resident in HBM) and on the CPU (bounded CPU-baseline samples).
Every generator returns a uint8 tensor of shape (n, 16): the reference's 16-byte Point records
(point-converter/src/point.rs:25-37: f32 x, y, z little-endian + r, g, b, a)."""
from __future__ import annotations

import torch


def _hash01(ix: torch.Tensor, iy: torch.Tensor, seed: int) -> torch.Tensor:
    """Integer lattice hash -> uniform [0, 1) float32 (splitmix-style, int64 arithmetic wraps)."""
    h = ix * 0x632BE59B + iy * 0x2545F491 + seed * 0x9E3779B1
    h = (h ^ (h >> 15)) * 0x2C1B3C6D
    h = (h ^ (h >> 12)) * 0x297A2D39
    h = h ^ (h >> 15)
    return (h & 0xFFFFFF).to(torch.float32) / 16777216.0


def _value_noise(x: torch.Tensor, y: torch.Tensor, wavelength: float, seed: int) -> torch.Tensor:
    u, v = x / wavelength, y / wavelength
    iu, iv = torch.floor(u), torch.floor(v)
    fu, fv = u - iu, v - iv
    iu, iv = iu.to(torch.int64), iv.to(torch.int64)
    su, sv = fu * fu * (3 - 2 * fu), fv * fv * (3 - 2 * fv)
    a, b = _hash01(iu, iv, seed), _hash01(iu + 1, iv, seed)
    c, d = _hash01(iu, iv + 1, seed), _hash01(iu + 1, iv + 1, seed)
    return (a + (b - a) * su) * (1 - sv) + (c + (d - c) * su) * sv


def _pack(xyz: torch.Tensor, gen: torch.Generator) -> torch.Tensor:
    n = xyz.shape[0]
    out = torch.empty((n, 16), dtype=torch.uint8, device=xyz.device)
    out[:, :12] = xyz.contiguous().view(torch.uint8).view(n, 12)
    rgb = torch.randint(0, 256, (n, 3), dtype=torch.uint8, device=xyz.device, generator=gen)
    out[:, 12:15] = rgb
    out[:, 15] = 255
    return out


def terrain(n: int, seed: int = 2, device="cpu", extent=(-2000.0, 2000.0, -2000.0, 2000.0), chunk=1 << 24,
            surface_seed: int | None = None) -> torch.Tensor:
    """C2: 2.5-D LiDAR-like terrain: x, y uniform in the extent, z = 3-octave value noise (amplitude 120 m,
    base wavelength 1 km) + Gaussian sensor noise sigma = 0.05 m.  Density for the named config:
    1e8 points over 4 km x 4 km = 6.25 points/m^2.  `surface_seed` (default: seed) picks the height field, `seed`
    the sample positions: slices of one big survey share the surface and differ in the positions."""
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    out = torch.empty((n, 16), dtype=torch.uint8, device=device)
    x0, x1, y0, y1 = extent
    for s in range(0, n, chunk):
        m = min(chunk, n - s)
        x = torch.rand(m, device=device, generator=gen) * (x1 - x0) + x0
        y = torch.rand(m, device=device, generator=gen) * (y1 - y0) + y0
        ss = seed if surface_seed is None else surface_seed
        z = (_value_noise(x, y, 1000.0, ss) + 0.5 * _value_noise(x, y, 500.0, ss + 1) +
             0.25 * _value_noise(x, y, 250.0, ss + 2)) * (120.0 / 1.75)
        z = z + torch.randn(m, device=device, generator=gen) * 0.05
        out[s:s + m] = _pack(torch.stack([x, y, z], dim=1).to(torch.float32), gen)
    return out


def uniform_cube(n: int, seed: int = 1, device="cpu") -> torch.Tensor:
    """C1: uniform points in the unit cube."""
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    return _pack(torch.rand((n, 3), device=device, generator=gen), gen)


def gaussian_blobs(n: int, seed: int = 4, device="cpu", blobs: int = 64, chunk=1 << 24,
                   sample_seed: int | None = None) -> torch.Tensor:
    """C4: Zipf-weighted isotropic Gaussian blobs, sigma log-uniform 1..100 m, centres in +-4 km.  `seed` fixes the
    blobs, `sample_seed` (default: continue the same stream) the samples."""
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    centres = torch.rand((blobs, 3), device=device, generator=gen) * torch.tensor([8000.0, 8000.0, 500.0], device=device) \
        - torch.tensor([4000.0, 4000.0, 0.0], device=device)
    sigma = 100.0 ** torch.rand(blobs, device=device, generator=gen)
    w = 1.0 / torch.arange(1, blobs + 1, device=device, dtype=torch.float32)
    if sample_seed is not None:
        gen.manual_seed(sample_seed)
    out = torch.empty((n, 16), dtype=torch.uint8, device=device)
    for s in range(0, n, chunk):
        m = min(chunk, n - s)
        k = torch.multinomial(w, m, replacement=True, generator=gen)
        xyz = centres[k] + torch.randn((m, 3), device=device, generator=gen) * sigma[k, None]
        out[s:s + m] = _pack(xyz.to(torch.float32), gen)
    return out


def city(n: int, seed: int = 5, device="cpu", extent: float = 10_000.0, boxes: int = 20_000, chunk=1 << 24,
         sample_seed: int | None = None) -> torch.Tensor:
    """C5: city-scale scan: `extent` x `extent` metres (10 km: ~100 level-0 cells), 60 % of the points on
    the ground (value noise +-2 m), 40 % on the faces of `boxes` axis-aligned boxes (footprint 10-60 m, height 5-120 m),
    sigma = 2 cm sensor noise.  `seed` fixes the city (boxes, ground), `sample_seed` (default: seed) the sample
    positions: slices of one big scan share the city and differ in the samples."""
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    bx = torch.rand((boxes, 2), device=device, generator=gen) * extent          # box corner (x, y)
    bs = 10.0 + 50.0 * torch.rand((boxes, 2), device=device, generator=gen)     # footprint
    bh = 5.0 + 115.0 * torch.rand(boxes, device=device, generator=gen) ** 3     # mostly low, a few towers
    if sample_seed is not None:
        gen.manual_seed(sample_seed)
    out = torch.empty((n, 16), dtype=torch.uint8, device=device)
    for s in range(0, n, chunk):
        m = min(chunk, n - s)
        u = torch.rand((m, 3), device=device, generator=gen)
        kind = torch.rand(m, device=device, generator=gen)
        # ground
        gx, gy = u[:, 0] * extent, u[:, 1] * extent
        gz = (_value_noise(gx, gy, 200.0, seed) - 0.5) * 4.0
        # box faces: roof (30 %) or one of the four walls
        k = torch.randint(0, boxes, (m,), device=device, generator=gen)
        face = torch.randint(0, 5, (m,), device=device, generator=gen)
        fx = bx[k, 0] + u[:, 0] * bs[k, 0]
        fy = bx[k, 1] + u[:, 1] * bs[k, 1]
        fz = u[:, 2] * bh[k]
        fx = torch.where(face == 1, bx[k, 0], torch.where(face == 2, bx[k, 0] + bs[k, 0], fx))
        fy = torch.where(face == 3, bx[k, 1], torch.where(face == 4, bx[k, 1] + bs[k, 1], fy))
        fz = torch.where(face == 0, bh[k], fz)
        on_box = kind < 0.4
        xyz = torch.stack([torch.where(on_box, fx, gx), torch.where(on_box, fy, gy), torch.where(on_box, fz, gz)], dim=1)
        xyz = xyz + torch.randn((m, 3), device=device, generator=gen) * 0.02
        out[s:s + m] = _pack(xyz.to(torch.float32), gen)
    return out
