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


def two_body_impact(n, seed=5, r_target=6.4, r_impactor=3.0):
    """Config 3 shape: a target and an impactor, each a Tillotson iron core
    (mat_id 100) inside a granite mantle (mat_id 101), on a jittered lattice in
    a NON-periodic box; internal units of examples/Planetary/EarthImpact
    (1e27 g, 1e8 cm, 1e8 cm/s).  Free surfaces, a density jump at the
    core-mantle boundary, and internal energies spread over the condensed,
    hybrid and expanded Tillotson branches.  Returns the usual dict plus
    "dim" (box size) and "h_max"."""
    rng = np.random.default_rng(seed)
    vol = 4.0 / 3.0 * np.pi * (r_target ** 3 + r_impactor ** 3)
    delta = (vol / n) ** (1.0 / 3.0)
    box = 4.0 * (r_target + r_impactor)
    c_t = np.array([0.40 * box, 0.5 * box, 0.5 * box])
    c_i = c_t + np.array([r_target + r_impactor + 2.0 * delta, 0.3 * r_target, 0.0])
    xs, mats, vs = [], [], []
    for centre, radius, vel in ((c_t, r_target, 0.0), (c_i, r_impactor, -0.01)):
        m = int(np.ceil(radius / delta)) + 1
        g = np.arange(-m, m + 1) * delta
        p = np.stack(np.meshgrid(g, g, g, indexing="ij"), axis=-1).reshape(-1, 3)
        p = p + rng.uniform(-0.5, 0.5, p.shape) * 0.2 * delta
        r = np.linalg.norm(p, axis=1)
        keep = r < radius
        p, r = p[keep], r[keep]
        xs.append(p + centre)
        mats.append(np.where(r < 0.55 * radius, 100, 101).astype(np.int32))
        v = np.zeros((len(p), 3), dtype=np.float32)
        v[:, 0] = vel
        vs.append(v)
    x = np.concatenate(xs)
    mat = np.concatenate(mats)
    v = np.concatenate(vs) + rng.uniform(-1e-4, 1e-4, (len(x), 3)).astype(np.float32)
    nn = len(x)
    rho0 = np.where(mat == 100, 7.8e-3, 2.68e-3)
    mass = (rho0 * delta ** 3 * rng.uniform(0.8, 1.25, nn)).astype(np.float32)
    u = np.exp(rng.uniform(np.log(1e-7), np.log(3e-5), nn)).astype(np.float32)
    h0 = np.float32(ETA * delta)
    ic = {"x": np.ascontiguousarray(x, dtype=np.float64), "v": np.ascontiguousarray(v, dtype=np.float32),
          "mass": mass, "h": np.full(nn, h0, dtype=np.float32), "u": u,
          "id": np.arange(1, nn + 1, dtype=np.int64), "mat_id": mat, "time_bin": None,
          "dim": (box, box, box), "h_max": float(1.3 * h0)}
    return ic


def top_grid(dim, h_max, max_top=12, stretch=1.1, gamma=1.825742):
    """cdim of the top-level grid the reference (and the library) builds: src/space_regrid.c:120-127."""
    dmax = max(dim)
    cell_min = 0.99 * dmax / max_top
    return tuple(int(np.floor(d / max(np.float32(h_max) * np.float32(gamma) * np.float32(stretch), cell_min)))
                 for d in dim)


def cell_block_random(cells, cdim, dim, per_cell, seed=42, h=None):
    """Config 4 shape, generated cell by cell: `per_cell` uniform-random particles in each
    top-level cell listed in `cells` (flat indices, x-major as the reference numbers them), each
    cell seeded by its own index, so any process generates the same particles for the same cell.
    ids are cell * 2^24 + k.  Returns the usual dict plus "cell" (top-level cell of each particle)."""
    cells = np.asarray(cells, dtype=np.int64)
    width = np.array([dim[k] / cdim[k] for k in range(3)])
    n = len(cells) * per_cell
    x = np.empty((n, 3), dtype=np.float64)
    v = np.empty((n, 3), dtype=np.float32)
    for q, c in enumerate(cells):
        rng = np.random.default_rng([seed, int(c)])
        i, j, k = c // (cdim[1] * cdim[2]), (c // cdim[2]) % cdim[1], c % cdim[2]
        lo = np.array([i, j, k]) * width
        u = rng.random((per_cell, 3))
        xx = lo + u * width
        # keep every particle strictly inside its cell after rounding
        xx = np.minimum(xx, np.nextafter(lo + width, lo))
        x[q * per_cell:(q + 1) * per_cell] = xx
        v[q * per_cell:(q + 1) * per_cell] = rng.uniform(-0.05, 0.05, (per_cell, 3))
    ntot_density = per_cell / float(np.prod(width))
    h0 = np.float32(ETA / ntot_density ** (1.0 / 3.0)) if h is None else np.float32(h)
    ic = {"x": x, "v": v, "mass": np.full(n, 1.0 / ntot_density, dtype=np.float32),
          "h": np.full(n, h0, dtype=np.float32), "u": np.ones(n, dtype=np.float32),
          "id": (np.repeat(cells, per_cell) << 24) + np.tile(np.arange(per_cell, dtype=np.int64), len(cells)),
          "mat_id": None, "time_bin": None, "cell": np.repeat(cells, per_cell)}
    return ic


def clump_and_strays(n_clump=600, n_stray=40, seed=3, h=0.03):
    """A clump in one corner of an open unit box and a few stray particles far from everything: with a small
    h_max the strays end the ghost at h_max without a neighbour (hydro_part_has_no_neighbours)."""
    rng = np.random.default_rng(seed)
    n = n_clump + n_stray
    x = np.concatenate([rng.random((n_clump, 3)) * 0.2 + 0.05, rng.random((n_stray, 3)) * 0.5 + 0.45])
    return {"x": np.ascontiguousarray(x, dtype=np.float64), "v": rng.uniform(-0.05, 0.05, (n, 3)).astype(np.float32),
            "mass": np.full(n, 1.0 / n, dtype=np.float32), "h": np.full(n, h, dtype=np.float32),
            "u": (1.0 + 0.5 * rng.random(n)).astype(np.float32), "id": np.arange(1, n + 1, dtype=np.int64),
            "mat_id": None, "time_bin": None}


def permute(ic, order):
    """Reorders every per-particle array with `order` (cell order)."""
    out = {}
    for k, a in ic.items():
        if isinstance(a, np.ndarray):
            out[k] = np.ascontiguousarray(a[order])
        elif a is None:
            out[k] = None
    return out
