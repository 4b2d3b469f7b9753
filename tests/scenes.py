"""Edge-case particle scenes for the parity tests (TEST INFRASTRUCTURE): inputs other than the reference's own dam break,
chosen to hit what the domain calls collisions, empties and boundaries.  All positions lie inside the Advect clamp box
[(-0.5+0.1)/n, 1+(-0.5-0.1)/n] (Simulation3D.cpp:227-230), which is where every particle is after one reference step."""
import numpy as np


def clamp_box(n):
    lo = np.float32((np.float32(-0.5) + np.float32(0.1)) / np.float32(n))
    hi = np.float32(np.float32(1.0) + (np.float32(-0.5) - np.float32(0.1)) / np.float32(n))
    return lo, hi


def _finish(pos, rng, n):
    lo, hi = clamp_box(n)
    pos = np.clip(pos.astype(np.float32), lo, hi)
    vel = rng.uniform(-1.5, 1.5, pos.shape).astype(np.float32)
    return np.ascontiguousarray(np.concatenate([pos, vel], axis=1), np.float32)


def crowded(n, seed=1):
    """Collisions: 1500 particles inside three neighbouring cells, with exact duplicates (distance ties -> lowest index wins,
    Simulation3D.cpp:296) and hundreds of particles per half-cell."""
    rng = np.random.default_rng(seed)
    centres = np.array([[5, 5, 5], [6, 5, 5], [5, 6, 6]], np.float32)
    pos = (centres[rng.integers(0, 3, 1500)] + rng.uniform(-0.5, 0.5, (1500, 3))) / n
    pos[100:140] = pos[60:100]          # duplicates at a higher index
    pos[700] = pos[0]                   # and of particle 0 (never propagated by the sweeps: index <= 0 is skipped, :242)
    return _finish(pos, rng, n)


def lattice(n, seed=2):
    """Boundaries: particles exactly on cell centres, exactly on half-cell faces (n*X = k + 0.5: roundf goes away from zero,
    :286-288) and on all six faces / corners of the clamp box."""
    rng = np.random.default_rng(seed)
    k = np.stack(np.meshgrid(np.arange(2, 10), np.arange(2, 8), np.arange(3, 9), indexing="ij"), -1).reshape(-1, 3).astype(np.float32)
    centres = k / n
    faces = (k[::3] + np.float32(0.5)) / n
    lo, hi = clamp_box(n)
    walls = rng.uniform(0.1, 0.9, (120, 3)).astype(np.float32)
    for i in range(120):
        walls[i, i % 3] = lo if (i // 3) % 2 == 0 else hi
    corners = np.array([[a, b, c] for a in (lo, hi) for b in (lo, hi) for c in (lo, hi)], np.float32)
    return _finish(np.concatenate([centres, faces, walls, corners]), rng, n)


def sparse(n, seed=3):
    """Empties: 300 particles scattered over the whole box -- almost every cell is empty, the level set and the layered
    extrapolation reach across the whole grid."""
    rng = np.random.default_rng(seed)
    lo, hi = clamp_box(n)
    return _finish(rng.uniform(lo, hi, (300, 3)), rng, n)


def minimal(n, seed=4):
    """The smallest particle count Simulate accepts: 273 (the reference reads particles[272], :198-199), in one corner."""
    rng = np.random.default_rng(seed)
    return _finish(rng.uniform(0.0, 0.3, (273, 3)), rng, n)


SCENES = {"crowded": crowded, "lattice": lattice, "sparse": sparse, "minimal": minimal}


def grids(n, seed=9):
    """Random MAC grids U, V, W in the reference layout."""
    rng = np.random.default_rng(seed)
    return tuple(rng.uniform(-0.5, 0.5, k).astype(np.float32) for k in ((n + 1) * n * n, n * (n + 1) * n, n * n * (n + 1)))
