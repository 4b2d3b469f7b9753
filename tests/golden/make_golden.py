"""Generates tests/golden/* from the UNMODIFIED reference (oracle/_ref/libfs3ref.so).

Run in the build container (needs /root/reference to build oracle/_ref):
    python tests/golden/make_golden.py
Outputs
  kat.json          known-answer table in the format of SURVEY.md Appendix C (cell hash, occupancy,
                    sums of U,V,W,phi, #phi<0) for 16^3 and 32^3 at steps 0,1,2,3,10 (+100 at 32^3)
  ref16_step{1,3}.npz  full reference state (particles, U, V, W, phi) for the 16^3 dam break
  ref24_stages.npz  per-stage reference outputs at 24^3, step 2 (state before, then after each stage)
"""
import json
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import RefSim, cellhash, build_oracle  # noqa: E402


def row(sim, n, step):
    p = sim.get_particles()
    u, v, w = sim.get_velocity()
    pd = (p[:, :3].astype(np.float32) * np.float32(n)).astype(np.float64)
    c = np.where(pd >= 0, np.floor(pd + 0.5), np.ceil(pd - 0.5)).astype(np.int64)
    key = c[:, 0] + n * (c[:, 1] + n * c[:, 2])
    cnt = np.bincount(key, minlength=n ** 3)
    r = dict(n=n, step=step, particles=int(p.shape[0]), cellhash=cellhash(p, n), max_per_cell=int(cnt.max()),
             occupied=int((cnt > 0).sum()), sumU=float(u.astype(np.float64).sum()),
             sumV=float(v.astype(np.float64).sum()), sumW=float(w.astype(np.float64).sum()))
    if step > 0:
        phi = sim.get_phi()
        r["sumPhi"] = float(phi.astype(np.float64).sum())
        r["nFluid"] = int((phi < 0).sum())
    return r


def main():
    build_oracle()
    rows = []
    for n, steps in ((16, (0, 1, 2, 3, 10)), (32, (0, 1, 2, 3, 10, 100))):
        sim = RefSim(n)
        done = 0
        for s in steps:
            while done < s:
                sim.step()
                done += 1
            rows.append(row(sim, n, s))
            if n == 16 and s in (1, 3):
                u, v, w = sim.get_velocity()
                np.savez_compressed(os.path.join(HERE, "ref16_step%d.npz" % s), particles=sim.get_particles(),
                                    u=u, v=v, w=w, phi=sim.get_phi())
        sim.close()
    json.dump(rows, open(os.path.join(HERE, "kat.json"), "w"), indent=1)

    # per-stage dump at 24^3
    n = 24
    sim = RefSim(n)
    sim.step(); sim.step()
    d = {}
    d["p0"] = sim.get_particles(); d["u0"], d["v0"], d["w0"] = sim.get_velocity(); d["phi0"] = sim.get_phi()
    sim.advect(); d["p_advect"] = sim.get_particles()
    sim.levelset(); d["phi_levelset"] = sim.get_phi()
    sim.p2g(); d["u_p2g"], d["v_p2g"], d["w_p2g"] = sim.get_velocity()
    sim.bodyforces(); sim.project(); d["u_proj"], d["v_proj"], d["w_proj"] = sim.get_velocity()
    np.savez_compressed(os.path.join(HERE, "ref24_stages.npz"), **d)
    for r in rows:
        print(r)


if __name__ == "__main__":
    main()
