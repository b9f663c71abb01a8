"""Seeded collocation sets generated on the device.

The reference samples with `pyDOE.lhs` on the host (S/helmholtz_2D.py:202) or a Python rejection
loop (S/gross_pitaevskii_2D.py:277-295).  At 16M-64M points that is the slowest part of a step, so the
benchmark / multi-GPU path draws a Latin-hypercube sample directly in HBM: per dimension a random
permutation of the n strata plus a uniform jitter inside each stratum.  The generators run on whatever device they
are given (the property tests use the CPU).  Data-parallel runs pass `seed + rank`: different streams per rank."""
import torch


def latin_hypercube(n, d, lo, hi, seed=1234, device="cuda", dtype=torch.float32):
    """[n,d] Latin-hypercube sample of the box [lo,hi]^d (lo/hi scalars or length-d sequences)."""
    gen = torch.Generator(device=device)
    gen.manual_seed(int(seed))
    cols = []
    for _ in range(d):
        perm = torch.randperm(n, generator=gen, device=device)
        cols.append((perm.to(torch.float64) + torch.rand(n, generator=gen, device=device, dtype=torch.float64)) / n)
    u = torch.stack(cols, dim=1)
    lo_t = torch.as_tensor(lo, dtype=torch.float64, device=device)
    hi_t = torch.as_tensor(hi, dtype=torch.float64, device=device)
    return (lo_t + (hi_t - lo_t) * u).to(dtype).contiguous()


def square_boundary(n, lo, hi, seed=1234, device="cuda", dtype=torch.float32):
    """n points on the perimeter of [lo,hi]^2 (n/4 per edge), for boundary_loss / loss_data."""
    gen = torch.Generator(device=device)
    gen.manual_seed(int(seed) + 7)
    k = max(n // 4, 1)
    t = lo + (hi - lo) * torch.rand(4, k, generator=gen, device=device, dtype=torch.float64)
    lo_c, hi_c = torch.full_like(t[0], lo), torch.full_like(t[0], hi)
    pts = torch.cat([torch.stack([t[0], lo_c], 1), torch.stack([t[1], hi_c], 1), torch.stack([lo_c, t[2]], 1),
                     torch.stack([hi_c, t[3]], 1)], dim=0)
    return pts.to(dtype).contiguous()


def disc_points(n, center=(0.0, 0.0), radius=1.0, seed=1234, device="cuda", dtype=torch.float32):
    """[n,2] collocation points in a disc, drawn the way S/gross_pitaevskii_2D.py:277-295 draws them: angle ~ U(0, 2 pi)
    and r ~ U(0, R) independently.  That is NOT uniform in area (the density goes like 1/r, points crowd the centre) -- the
    reference trains on exactly this distribution, so it is reproduced, without the 10^4-iteration python loop."""
    gen = torch.Generator(device=device)
    gen.manual_seed(int(seed))
    theta = 2.0 * torch.pi * torch.rand(n, generator=gen, device=device, dtype=torch.float64)
    r = float(radius) * torch.rand(n, generator=gen, device=device, dtype=torch.float64)
    pts = torch.stack([center[0] + r * torch.cos(theta), center[1] + r * torch.sin(theta)], dim=1)
    return pts.to(dtype).contiguous()


def disc_boundary(n, center=(0.0, 0.0), radius=1.0, device="cuda", dtype=torch.float32):
    """n boundary points on the circle, theta = linspace(0, 2 pi, n) (first = last, as in :278-281)."""
    theta = torch.linspace(0.0, 2.0 * torch.pi, n, device=device, dtype=torch.float64)
    pts = torch.stack([center[0] + radius * torch.cos(theta), center[1] + radius * torch.sin(theta)], dim=1)
    return pts.to(dtype).contiguous()
