"""Synthetic initial conditions of the shapes BASELINE.json:configs names
(SURVEY.md section 8(d)).  numpy only; deterministic in the seed."""
import numpy as np

ETA = 1.2348  # SPH:resolution_eta, examples/HydroTests/SedovBlast_3D/sedov.yml:29


def _finish(x, v, u, rho0=1.0, box=1.0, mat_id=None, h=None):
    n = len(x)
    ic = {
        "x": np.ascontiguousarray(x, dtype=np.float64),
        "v": np.ascontiguousarray(v, dtype=np.float32),
        "mass": np.full(n, rho0 * box ** 3 / n, dtype=np.float32),
        "h": (np.full(n, ETA * box / round(n ** (1 / 3)), dtype=np.float32) if h is None
              else np.ascontiguousarray(h, dtype=np.float32)),
        "u": np.ascontiguousarray(u, dtype=np.float32),
        "id": np.arange(1, n + 1, dtype=np.int64),
        "mat_id": mat_id,
        "time_bin": None,
    }
    return ic


def lattice(nside, jitter=0.1, seed=1234, box=1.0, sedov=True):
    """Config 1 shape: nside^3 simple-cubic lattice in a periodic unit box,
    jittered by `jitter` inter-particle spacings, Sedov-like energy spike."""
    rng = np.random.default_rng(seed)
    g = (np.arange(nside) + 0.5) / nside
    x = np.stack(np.meshgrid(g, g, g, indexing="ij"), axis=-1).reshape(-1, 3) * box
    if jitter > 0:
        x = x + rng.uniform(-0.5, 0.5, x.shape) * jitter * box / nside
    x = np.mod(x, box)
    x[x >= box] = 0.0
    n = len(x)
    u = np.full(n, 1e-6 / ((5.0 / 3.0 - 1.0) * 1.0), dtype=np.float32)
    if sedov:
        r2 = ((x - 0.5 * box) ** 2).sum(1)
        centre = np.argsort(r2)[:15]
        u[centre] = 1.0 / (15 * (box ** 3 / n))
    v = np.zeros((n, 3), dtype=np.float32)
    return _finish(x, v, u, box=box)


def uniform_random(n, seed=42, box=1.0):
    """Config 2/4 shape: uniform-random positions in a periodic unit box,
    equal masses, v ~ U(-0.05, 0.05), u = 1."""
    rng = np.random.default_rng(seed)
    x = rng.random((n, 3)) * box
    x[x >= box] = 0.0
    v = rng.uniform(-0.05, 0.05, (n, 3)).astype(np.float32)
    u = np.ones(n, dtype=np.float32)
    ic = _finish(x, v, u, box=box)
    ic["h"][:] = ETA * box / n ** (1.0 / 3.0)
    return ic


def clustered(n, nclumps=64, seed=7, box=1.0, frac_background=0.3):
    """Config 5 shape: Plummer-like clumps over a uniform background; wide h range, deep tree."""
    rng = np.random.default_rng(seed)
    nb = int(n * frac_background)
    nc = n - nb
    xb = rng.random((nb, 3)) * box
    centres = rng.random((nclumps, 3)) * box
    scale = box * 0.01 * (1 + 4 * rng.random(nclumps))
    which = rng.integers(0, nclumps, nc)
    # Plummer radius sampling
    m = rng.random(nc) * 0.95 + 0.01
    r = scale[which] / np.sqrt(m ** (-2.0 / 3.0) - 1.0)
    d = rng.normal(size=(nc, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    xc = centres[which] + d * r[:, None]
    x = np.mod(np.concatenate([xb, xc]), box)
    x[x >= box] = 0.0
    v = rng.uniform(-0.05, 0.05, (n, 3)).astype(np.float32)
    u = np.ones(n, dtype=np.float32)
    ic = _finish(x, v, u, box=box)
    # local-density based first guess of h (k-th neighbour is too slow: use clump scale)
    hb = np.full(nb, ETA * box / max(nb, 1) ** (1 / 3), dtype=np.float32)
    per_clump = np.maximum(np.bincount(which, minlength=nclumps), 1)
    hc = (ETA * 2.0 * np.maximum(r, 0.3 * scale[which]) / per_clump[which] ** (1 / 3)).astype(np.float32)
    ic["h"] = np.concatenate([hb, np.minimum(hc, hb[0] if nb else 1.0)]).astype(np.float32)
    return ic


def permute(ic, order):
    """Reorders every per-particle array with `order` (cell order)."""
    out = {}
    for k, a in ic.items():
        out[k] = None if a is None else np.ascontiguousarray(a[order])
    return out
