"""On-disk state format (include/fs3.h fs3_state_*).

CPU part: the host-only file functions of the C-ABI library -- round trip of an oracle state, layout of the
bytes on disk, and loud failures on every kind of damage.  GPU part (`-m gpu`): save -> destroy -> create -> load
continues bit-identically, for a single handle and for a z-slab sharded group (one file per rank).
"""
import os
import struct
import numpy as np
import pytest

from oracle_lib import Oracle


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def oracle_state(n, steps):
    ora = Oracle(n)
    for _ in range(steps):
        ora.step()
    return ora, ora.get_particles(), ora.get_velocity(), ora.get_phi()


def py_checksum(chunks):
    """Independent restatement of fs3_state_checksum: FNV-1a over little-endian 64-bit words, chained over sections."""
    h = 0
    for c in chunks:
        b = np.ascontiguousarray(c).tobytes()
        if not b:
            continue
        b += b"\0" * (-len(b) % 8)
        h = h or 0xcbf29ce484222325
        for (w,) in struct.iter_unpack("<Q", b):
            h = ((h ^ w) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
    return h


def test_round_trip_and_layout(tmp_path):
    from fluidsimulation_b200 import state_file_write, state_file_read, state_file_info
    n = 16
    _, p, (u, v, w), phi = oracle_state(n, 2)
    path = str(tmp_path / "a.fs3")
    state_file_write(path, n, p, u, v, w, phi, steps=2)
    assert not os.path.exists(path + ".tmp")
    info = state_file_info(path)
    assert (info.version, info.nx, info.ny, info.nz, info.cells_per_meter) == (1, n, n, n, float(n))
    assert (info.num_particles, info.total_particles, info.steps, info.rank, info.world, info.has_ids) == (p.shape[0], p.shape[0], 2, 0, 1, 0)
    assert info.params.sor_iterations == 100 and info.params.rho == 1000.0 and info.params.force_dt_0_01 == 1   # fs3_default_params
    info2, p2, ids2, u2, v2, w2, phi2 = state_file_read(path)
    assert ids2 is None and info2.checksum == info.checksum
    for a, b in ((p, p2), (u, u2), (v, v2), (w, w2), (phi, phi2)):
        assert np.array_equal(bits(a), bits(b))
    # bytes on disk: 256-byte header, then the sections back to back, little-endian
    raw = open(path, "rb").read()
    assert raw[:8] == b"FS3STATE" and struct.unpack_from("<II", raw, 8) == (1, 256)
    assert struct.unpack_from("<iiif", raw, 16) == (n, n, n, float(n))
    assert struct.unpack_from("<qqq", raw, 32) == (p.shape[0], p.shape[0], 2)
    payload = p.tobytes() + u.tobytes() + v.tobytes() + w.tobytes() + phi.tobytes()
    assert raw[256:] == payload and info.payload_bytes == len(payload)
    assert info.checksum == py_checksum([p, u, v, w, phi])


def test_checksum_matches_restatement():
    from fluidsimulation_b200 import state_checksum
    rng = np.random.default_rng(0)
    for nbytes in (0, 1, 7, 8, 9, 4096, 4099):
        a = rng.integers(0, 256, nbytes, dtype=np.uint8)
        if nbytes:
            assert state_checksum(a) == py_checksum([a])
    a, b = rng.integers(0, 256, 36, dtype=np.uint8), rng.integers(0, 256, 20, dtype=np.uint8)
    assert state_checksum(b, state_checksum(a)) == py_checksum([a, b])


def test_sharded_file_with_ids(tmp_path):
    from fluidsimulation_b200 import state_file_write, state_file_read
    n = 16
    _, p, (u, v, w), phi = oracle_state(n, 0)
    mine = np.flatnonzero(np.round(p[:, 2] * n) < n // 2).astype(np.int32)
    path = str(tmp_path / "rank0.fs3")
    state_file_write(path, n, p[mine], u, v, w, phi, ids=mine, rank=0, world=2, total_particles=p.shape[0],
                     rank_particles=[mine.size, p.shape[0] - mine.size])
    info, p2, ids2, *_ = state_file_read(path)
    assert (info.rank, info.world, info.has_ids, info.total_particles) == (0, 2, 1, p.shape[0])
    assert list(info.rank_particles)[:3] == [mine.size, p.shape[0] - mine.size, 0]
    assert np.array_equal(ids2, mine) and np.array_equal(bits(p2), bits(p[mine]))


def test_damage_is_reported(tmp_path):
    from fluidsimulation_b200 import state_file_write, state_file_read, state_file_info, Fs3Error
    n = 16
    _, p, (u, v, w), phi = oracle_state(n, 0)
    good = str(tmp_path / "good.fs3")
    state_file_write(good, n, p, u, v, w, phi)
    raw = bytearray(open(good, "rb").read())

    def variant(name, data):
        q = str(tmp_path / name)
        open(q, "wb").write(bytes(data))
        return q

    with pytest.raises(Fs3Error, match="cannot open"):
        state_file_info(str(tmp_path / "missing.fs3"))
    with pytest.raises(Fs3Error, match="shorter than"):
        state_file_info(variant("short.fs3", raw[:100]))
    bad = bytearray(raw); bad[0:8] = b"NOTSTATE"
    with pytest.raises(Fs3Error, match="magic"):
        state_file_info(variant("magic.fs3", bad))
    bad = bytearray(raw); struct.pack_into("<I", bad, 8, 99)
    with pytest.raises(Fs3Error, match="version"):
        state_file_info(variant("version.fs3", bad))
    bad = bytearray(raw); struct.pack_into("<q", bad, 32, p.shape[0] + 1)
    with pytest.raises(Fs3Error, match="payload size"):
        state_file_info(variant("count.fs3", bad))
    with pytest.raises(Fs3Error, match="truncated"):
        state_file_read(variant("trunc.fs3", raw[:-4]))
    bad = bytearray(raw); bad[256 + 1000] ^= 0x40
    with pytest.raises(Fs3Error, match="checksum"):
        state_file_read(variant("flip.fs3", bad))
    # an unwritable target leaves nothing behind
    with pytest.raises(Fs3Error, match="cannot open"):
        state_file_write(str(tmp_path / "no_such_dir" / "x.fs3"), n, p, u, v, w, phi)


# ---- GPU: resume is bit-identical --------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("n", [16, 32])
def test_save_load_resume_bitexact(tmp_path, n):
    from fluidsimulation_b200 import FluidSim3, state_file_read
    path = str(tmp_path / "s.fs3")
    a = FluidSim3(n, n, n, float(n))
    a.set_params(sor_iterations=40)
    for _ in range(3):
        a.Simulate(0.01)
    a.save_state(path)
    assert a.step_count == 3
    saved = a.get_particles()
    for _ in range(2):
        a.Simulate(0.01)
    want_p, want_v, want_phi = a.get_particles(), a.get_velocity(), a.get_phi()
    a.close()
    # the file holds exactly what the handle held
    info, p, ids, u, v, w, phi = state_file_read(path)
    assert info.steps == 3 and info.params.sor_iterations == 40 and ids is None
    assert np.array_equal(bits(p), bits(saved))
    b = FluidSim3(n, n, n, float(n))
    b.load_state(path)
    assert b.step_count == 3 and b.get_params().sor_iterations == 40
    assert np.array_equal(bits(b.get_particles()), bits(saved))
    for _ in range(2):
        b.Simulate(0.01)
    assert b.step_count == 5
    assert np.array_equal(bits(b.get_particles()), bits(want_p)), "resumed run diverged"
    for x, y in zip(b.get_velocity(), want_v):
        assert np.array_equal(bits(x), bits(y))
    assert np.array_equal(bits(b.get_phi()), bits(want_phi))


@pytest.mark.gpu
def test_oracle_state_file_drives_the_gpu(tmp_path):
    """A file written on the host from the oracle's state loads into the GPU sim; one step then matches the oracle."""
    from fluidsimulation_b200 import FluidSim3, state_file_write
    n = 24
    ora, p, (u, v, w), phi = oracle_state(n, 2)
    path = str(tmp_path / "o.fs3")
    state_file_write(path, n, p, u, v, w, phi, steps=2)
    sim = FluidSim3(n, n, n, float(n))
    sim.load_state(path)
    sim.Simulate(0.01); ora.step()
    sim.sync()
    assert np.array_equal(sim.get_cell_indices(), ora.get_cell_indices())
    assert np.array_equal(bits(sim.get_phi()), bits(ora.get_phi()))
    assert float(np.abs(sim.get_particles() - ora.get_particles()).max()) < 1e-5   # north_star tolerance (P2G sum order)


@pytest.mark.gpu
def test_load_refuses_other_shapes(tmp_path):
    from fluidsimulation_b200 import FluidSim3, Fs3Error
    path = str(tmp_path / "s16.fs3")
    a = FluidSim3(16, 16, 16, 16.0)
    a.save_state(path)
    b = FluidSim3(24, 24, 24, 24.0)
    with pytest.raises(Fs3Error, match="different grid"):
        b.load_state(path)
    with pytest.raises(Fs3Error, match="cannot open"):
        b.load_state(str(tmp_path / "nope.fs3"))


@pytest.mark.gpu
def test_sharded_group_save_load_resume_bitexact(tmp_path):
    from fluidsimulation_b200 import FluidSim3Group
    n, world = 32, 2
    g = FluidSim3Group(n, world)
    for _ in range(2):
        g.Simulate(0.01)
    paths = [str(tmp_path / ("rank%d.fs3" % r)) for r in range(world)]
    for r, path in zip(g.ranks, paths):
        r.save_state(path)
    for _ in range(2):
        g.Simulate(0.01)
    g.sync()
    want = g.get_particles()
    g.close()
    h = FluidSim3Group(n, world)
    for r, path in zip(h.ranks, paths):
        r.load_state(path)
    for _ in range(2):
        h.Simulate(0.01)
    h.sync()
    assert np.array_equal(bits(h.get_particles()), bits(want)), "resumed sharded run diverged"


# ---- C++ host tool over the same ABI (examples/state_info.cpp) and the per-rank files of a sharded run (gloo, world 2) ----
def build_tool():
    import subprocess
    from fluidsimulation_b200 import build
    build.build()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe, src = os.path.join(root, "examples", "state_info"), os.path.join(root, "examples", "state_info.cpp")
    deps = [src, os.path.join(root, "include", "fs3.h")]
    if (not os.path.exists(exe)) or any(os.path.getmtime(d) > os.path.getmtime(exe) for d in deps):
        lib = os.path.join(root, "fluidsimulation_b200")
        subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-I" + os.path.join(root, "include"), src, "-L" + lib, "-lfluidsim3_b200",
                               "-Wl,-rpath," + lib, "-Wl,-rpath,$ORIGIN/../fluidsimulation_b200", "-o", exe])
    return exe


def test_cpp_state_tool(tmp_path):
    import subprocess
    from fluidsimulation_b200 import state_file_write
    exe = build_tool()
    n = 16
    _, p, (u, v, w), phi = oracle_state(n, 1)
    path, tsv = str(tmp_path / "t.fs3"), str(tmp_path / "slice.tsv")
    state_file_write(path, n, p, u, v, w, phi, steps=1)
    r = subprocess.run([exe, path, "phi", "5", tsv], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "grid 16x16x16" in r.stdout and "steps 1" in r.stdout and "OK" in r.stdout
    l2_phi = float(r.stdout.split("phi ")[-1].split()[0])
    assert abs(l2_phi - float(np.sqrt((phi.astype(np.float64) ** 2).sum()))) <= 1e-6 * l2_phi
    # the text slice round-trips fp32 exactly (%.9g), unlike the reference's 6-digit dumps
    sl = np.loadtxt(tsv, dtype=np.float64).astype(np.float32)
    assert np.array_equal(bits(sl), bits(phi.reshape(n, n, n)[:, 5, :]))
    raw = bytearray(open(path, "rb").read()); raw[300] ^= 1
    bad = str(tmp_path / "bad.fs3"); open(bad, "wb").write(bytes(raw))
    r = subprocess.run([exe, bad], capture_output=True, text=True)
    assert r.returncode == 1 and "checksum" in r.stderr


def test_per_rank_files_two_rank_gloo(tmp_path):
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29537", os.path.join(root, "tests", "state_shard_worker.py"), str(tmp_path)],
                       capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "state shard worker ok" in r.stdout
