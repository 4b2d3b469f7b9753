"""GPU worker for tests/test_levelset_lines.py (own process, ctypes + numpy only): the one-layer extrapolation
(fs3_params.extrapolation_mode = 1) bit-exact to its CPU restatement on arbitrary grids through the fs3_extrapolate hook,
and a whole step with both GPU-formulation modes switched on.  Prints GEX_WORKER_OK."""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def main():
    from fluidsimulation_b200 import FluidSim3
    from oracle_lib import gls_extrapolate, gls_levelset
    n = 16
    sim = FluidSim3(n, n, n, float(n))
    sim.set_params(extrapolation_mode=1)
    assert sim.get_params().extrapolation_mode == 1
    rng = np.random.default_rng(11)
    for dims, density in (((9, 8, 8), 0.05), ((8, 9, 8), 0.3), ((8, 8, 9), 0.0), ((5, 7, 6), 0.9), ((17, 16, 16), 0.2)):
        nn = dims[0] * dims[1] * dims[2]
        src = rng.standard_normal(nn).astype(np.float32)
        valid = (rng.random(nn) < density).astype(np.uint8)
        got = sim.extrapolate(src, valid, *dims)
        want = gls_extrapolate(src, valid, *dims)
        assert np.array_equal(bits(got), bits(want)), "dims %s density %g: %d nodes differ" % (dims, density, int((bits(got) != bits(want)).sum()))
    # whole steps with both GPU-formulation modes on: finite, phi is the line-sweep level set of the advected particles
    for m in (24, 32):
        s2 = FluidSim3(m, m, m, float(m))
        s2.set_params(levelset_mode=1, extrapolation_mode=1)
        for _ in range(3):
            s2.Simulate(0.01)
        s2.sync()
        p = s2.get_particles()
        assert np.isfinite(p).all() and all(np.isfinite(a).all() for a in s2.get_velocity())
        want_phi, _ = gls_levelset(p, m)
        assert np.array_equal(bits(s2.get_phi()), bits(want_phi))
        s2.close()
    sim.close()
    # z-slab sharded (2 ranks in lockstep on this GPU): the one-layer pass runs replicated on the gathered grids, bit-identical
    from fluidsimulation_b200 import FluidSim3Group
    one, grp = FluidSim3(16, 16, 16, 16.0), FluidSim3Group(16, 2)
    one.set_params(extrapolation_mode=1)
    grp.set_params(extrapolation_mode=1)
    for _ in range(3):
        one.Simulate(0.01)
        grp.Simulate(0.01)
    one.sync()
    grp.sync()
    assert np.array_equal(bits(grp.get_particles()), bits(one.get_particles())), "sharded one-layer extrapolation differs from single GPU"
    grp.close()
    one.close()
    print("GEX_WORKER_OK")


if __name__ == "__main__":
    main()
