"""The C++ host side (include/FluidSim3Compat.hpp + examples/dam_break.cpp): the reference's class surface
over the C ABI.  CPU: it compiles, links and fails loudly without a GPU.  GPU: its m_particles after N
frames match the oracle's."""
import os
import subprocess
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "examples", "dam_break")


def build_example():
    from fluidsimulation_b200 import build
    build.build()
    src = os.path.join(ROOT, "examples", "dam_break.cpp")
    deps = [src, os.path.join(ROOT, "include", "FluidSim3Compat.hpp"), os.path.join(ROOT, "include", "fs3.h")]
    if (not os.path.exists(EXE)) or any(os.path.getmtime(d) > os.path.getmtime(EXE) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), src,
                               "-L" + os.path.join(ROOT, "fluidsimulation_b200"), "-lfluidsim3_b200",
                               "-Wl,-rpath," + os.path.join(ROOT, "fluidsimulation_b200"),
                               "-Wl,-rpath,$ORIGIN/../fluidsimulation_b200", "-o", EXE])
    return EXE


def test_cpp_host_builds_and_has_no_cpu_fallback():
    import torch
    exe = build_example()
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = subprocess.run([exe, "16", "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "fs3_create" in r.stderr


@pytest.mark.gpu
def test_cpp_host_matches_oracle(tmp_path):
    from oracle_lib import Oracle
    exe = build_example()
    n, frames = 16, 3
    dump = str(tmp_path / "p.bin")
    r = subprocess.run([exe, str(n), str(frames), dump], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    p = np.fromfile(dump, np.float32).reshape(-1, 6)
    ora = Oracle(n)
    for _ in range(frames):
        ora.step(0.01)
    q = ora.get_particles()
    assert p.shape == q.shape
    assert float(np.abs(p - q).max()) <= 1e-5 * max(1.0, float(np.abs(q).max()))
