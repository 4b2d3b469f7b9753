"""Generates tests/golden/gpuform16.npz: outputs of the section-8(f) restatements (oracle/fs3_consumer_oracle.c,
oracle/fs3_gpu_semantics_oracle.c) on the committed reference state ref16_step1.npz.

These restate D3D11 shaders that cannot be run here (parity against the reference: UNPINNED), so this fixture does not pin
them to the reference; it freezes the deterministic reading the CUDA kernels are held to, so that a later edit of a
restatement shows up as a diff of this file instead of passing silently.
    python tests/golden/make_golden_8f.py
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import blur_phi, sample_phi, bin_particles, gls_levelset, gls_extrapolate  # noqa: E402


def inputs():
    n = 16
    ref = np.load(os.path.join(HERE, "ref16_step1.npz"))
    p, phi, u = ref["particles"], ref["phi"], ref["u"]
    rng = np.random.default_rng(816)
    pts = rng.uniform(-0.1, 1.1, (512, 3)).astype(np.float32)
    valid = (rng.random(u.size) < 0.15).astype(np.uint8)
    return n, p, phi, u, pts, valid


def compute():
    n, p, phi, u, pts, valid = inputs()
    blurred = blur_phi(phi, n)
    binned, offsets = bin_particles(p, n)
    gphi, gclosest = gls_levelset(p, n)
    return dict(blurred=blurred, samples=sample_phi(phi, n, pts), samples_blurred=sample_phi(blurred, n, pts), offsets=offsets,
                binned_first=binned[:64].copy(), binned_colsum=np.ascontiguousarray(binned).view(np.uint32).sum(0, dtype=np.uint64),
                gls_phi=gphi, gls_closest=gclosest, one_layer_U=gls_extrapolate(u, valid, n + 1, n, n))


if __name__ == "__main__":
    np.savez_compressed(os.path.join(HERE, "gpuform16.npz"), **compute())
    print("wrote gpuform16.npz")
