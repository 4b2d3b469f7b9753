"""GPU worker for tests/test_levelset_lines.py (runs in its own process: a fault in a new kernel must not poison the CUDA
context of the other tests).  ctypes + numpy only.  Prints GLS_WORKER_OK when the line-sweep level set
(fs3_params.levelset_mode = 1) is bit-exact to its CPU restatement on every case."""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def check(tag, sim, n, gls_levelset):
    p = sim.get_particles()
    sim.levelset()
    phi, closest = sim.get_phi(), sim.get_closest()
    want_phi, want_closest = gls_levelset(p, n)
    assert np.array_equal(bits(phi), bits(want_phi)), "%s: phi differs in %d cells" % (tag, int((bits(phi) != bits(want_phi)).sum()))
    assert np.array_equal(closest, want_closest), "%s: closest differs in %d cells" % (tag, int((closest != want_closest).sum()))
    assert np.array_equal(bits(sim.get_particles()), bits(p)), tag + ": the level set must not move particles"


def main():
    from fluidsimulation_b200 import FluidSim3
    from oracle_lib import gls_levelset
    from scenes import SCENES, grids
    sizes = [int(a) for a in sys.argv[1:]] or [16, 24]
    for n in sizes:
        sim = FluidSim3(n, n, n, float(n))
        sim.set_params(levelset_mode=1)
        assert sim.get_params().levelset_mode == 1
        check("dam break %d^3, reset state" % n, sim, n, gls_levelset)
        # a full step in this mode: phi at the end of the step is the level set of the advected particles
        sim.Simulate(0.01)
        sim.sync()
        p = sim.get_particles()
        assert np.isfinite(p).all()
        want_phi, want_closest = gls_levelset(p, n)
        assert np.array_equal(bits(sim.get_phi()), bits(want_phi)), "step %d^3: phi differs" % n
        assert np.array_equal(sim.get_closest(), want_closest), "step %d^3: closest differs" % n
        sim.Simulate(0.01)
        check("dam break %d^3, after 2 steps" % n, sim, n, gls_levelset)
        if n == 16:
            for name in sorted(SCENES):
                sim.set_particles(SCENES[name](n))
                sim.set_velocity(*grids(n))
                check("scene %s" % name, sim, n, gls_levelset)
        # back to the default mode: the parity path is untouched
        sim.set_params(levelset_mode=0)
        sim.Simulate(0.01)
        sim.sync()
        sim.close()
    print("GLS_WORKER_OK")


if __name__ == "__main__":
    main()
