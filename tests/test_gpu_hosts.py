"""The drop-in host programs (hosts/*.cpp, C++ through the C ABI) write the same files as the reference
programs: SHA-256 of the text output against the fingerprints BASELINE.md section 3 recorded from the
unmodified reference (binary.cpp, mrt.cpp, another_mrt.cpp), and the .vts of the QuarterChannel drop-in
parsed back and compared with the golden field samples."""
import hashlib
import os
import struct
import subprocess

import numpy as np
import pytest

import golden_lib as gl
from conftest import ROOT

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "hosts", "bin")


def _sha16(path):
    return hashlib.sha256(open(path, "rb").read()).hexdigest()[:16]


def _run(prog, args, cwd, env=None):
    exe = os.path.join(BIN, prog)
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", os.path.join(ROOT, "hosts")], check=True, capture_output=True)
    r = subprocess.run([exe] + args, cwd=cwd, capture_output=True, text=True, env=dict(os.environ, **(env or {})), timeout=900)
    assert r.returncode == 0, r.stderr[-500:]
    return r.stdout


def test_binary_program_files(tmp_path):
    out = _run("binary", [], str(tmp_path))
    assert "Time is 20000" in out
    assert _sha16(tmp_path / "phase020000.dat") == "97765059e339cfbc"
    assert _sha16(tmp_path / "velocityx020000.dat") == "adfd2c25b9864189"
    assert _sha16(tmp_path / "velocityy020000.dat") == "db6ef2429f6716b2"
    assert _sha16(tmp_path / "phase000100.dat") == "836c201b829113f9"
    assert len([f for f in os.listdir(tmp_path) if f.endswith(".dat")]) == 603


def test_mrt_program_files(tmp_path):
    os.mkdir(tmp_path / "tmp")
    out = _run("mrt", ["10", "10"], str(tmp_path))
    assert out.startswith("Omega=1\nOmega_bulk=1\n0\n")
    assert _sha16(tmp_path / "tmp" / "phase000800.dat") == "07ca3fdcfdd7cb4c"
    assert _sha16(tmp_path / "tmp" / "proba000800.dat") == "2070cdc933cf2c1d"


def test_another_mrt_program_files(tmp_path):
    os.mkdir(tmp_path / "tmp")
    _run("another_mrt", [], str(tmp_path))
    assert _sha16(tmp_path / "tmp" / "phase000500.dat") == "efa99204039552d8"
    assert _sha16(tmp_path / "tmp" / "proba000500.dat") == "45c1dd18a5ef069e"


def test_shanchen_program_runs(tmp_path):
    """Text output at 6 digits: compared value by value with the golden (shared-exp) rho of step 1000."""
    out = _run("shanchen", ["1000"], str(tmp_path))
    assert "Iteration is 1000" in out
    rows = [ln.split() for ln in open(tmp_path / "density01000.dat").read().splitlines() if ln.strip()]
    got = np.array(rows, dtype=np.float64).ravel()
    ref, ent = gl.samples("shanchen_sharedexp_t1000", "rho")
    assert ent.get("whole")
    want = np.array([float("%.6g" % v) for v in ref])
    assert np.array_equal(got, want)


def _read_vts(path):
    raw = open(path, "rb").read()
    head, rest = raw.split(b"<AppendedData encoding=\"raw\">", 1)
    data = rest[rest.index(b"_") + 1:]
    import re
    arrays = {}
    for m in re.finditer(rb'<DataArray type="(\w+)" Name="(\w+)"[^>]*offset="(\d+)"', head):
        typ, name, off = m.group(1).decode(), m.group(2).decode(), int(m.group(3))
        nbytes = struct.unpack("<Q", data[off:off + 8])[0]
        arrays[name] = np.frombuffer(data[off + 8:off + 8 + nbytes], dtype=np.float64 if typ == "Float64" else np.float32)
    return arrays


def test_microchannel_quarter_vts(tmp_path):
    out = _run("microchannel_quarter", [], str(tmp_path), env={"LBM_NUMTIME": "101"})
    assert "Output is done on the step 100" in out
    a = _read_vts(tmp_path / "phase00100.vts")
    assert list(a) == ["Velocity", "Phase", "Density", "Points"]
    vel = a["Velocity"].reshape(-1, 3)
    for name, arr in (("phase", a["Phase"]), ("rho", a["Density"]), ("ux", vel[:, 0]), ("uy", vel[:, 1]), ("uz", vel[:, 2])):
        assert gl.check("quarter_t100", name, np.ascontiguousarray(arr))[0] == 0, name
    assert a["Points"].reshape(-1, 3)[28].tolist() == [1.0, 1.0, 0.0]


def test_cylinder_program_files(tmp_path):
    """hosts/cylinder.cpp: the complete 64001-iteration run; fingerprints taken from the unmodified
    DiffusionCylinder/cylinder.cpp built with g++ -O2 -ffp-contract=off and run in an empty directory
    holding ./antibb (321 files)."""
    os.mkdir(tmp_path / "antibb")
    out = _run("cylinder", [], str(tmp_path))
    assert hashlib.sha256(out.encode()).hexdigest()[:16] == "6363b74654d9bd53"
    assert _sha16(tmp_path / "antibb" / "density0000000.dat") == "57cf3c67cd2af58f"
    assert _sha16(tmp_path / "antibb" / "density0000200.dat") == "3d0afc83ee2ff0a7"
    assert _sha16(tmp_path / "antibb" / "density0064000.dat") == "b719d0f7297f1ed1"
    assert len(os.listdir(tmp_path / "antibb")) == 321


def test_binary_shape_program_files(tmp_path):
    """hosts/binary_shape.cpp: the complete 20001-iteration run; fingerprints taken from the unmodified
    CurvedSolid/binary_shape.cpp built with g++ -O2 -ffp-contract=off and run in an empty directory
    (stdout with all 201 'Mass is' lines, geometry.dat, phase.dat, the single '.dat' the three field
    writers share, 101 legacy .vtk files)."""
    out = _run("binary_shape", [], str(tmp_path))
    assert hashlib.sha256(out.encode()).hexdigest()[:16] == "2b98d4974a4ac78c"
    assert _sha16(tmp_path / "geometry.dat") == "029458e9dfa10ff8"
    assert _sha16(tmp_path / "phase.dat") == "0ba98c3844c89b3e"
    assert _sha16(tmp_path / ".dat") == "d4cabfac3c757cf6"
    assert _sha16(tmp_path / "vtk0000000.vtk") == "69194701a6e5292e"
    assert _sha16(tmp_path / "vtk0000200.vtk") == "a6ec05415407c75b"
    assert _sha16(tmp_path / "vtk0020000.vtk") == "50ccb783563608e9"
    assert len(os.listdir(tmp_path)) == 104


# SHA-256 prefixes of the files the REFERENCE Solver's own text writers (solver.hpp:578-787, compiled into
# oracle/_ref/libref_quarter.so, ref3d_write_text) produce for the run hosts/solver_selftest.cpp makes:
# 14x11x16, phase_gradient 0.02, bubble width 4, 7 steps.
REFERENCE_TEXT_FILES = {
    "density_whole.dat": "fdc70e5cc6234728", "density_xy.dat": "00cf984c11a94a5e", "density_xz.dat": "b269441ccf0353a7",
    "density_yz.dat": "f525e1b38910c6c4", "phase_legacy.vtk": "dc4becf80d4993dc", "phase_whole.dat": "ad3ea1c91c740252",
    "phase_xy.dat": "777ce63a15d1b91d", "phase_xz.dat": "ec99e2896df95d9a", "phase_yz.dat": "8aecd41105084859",
    "vel_whole.dat": "73ee85e965e22489", "vel_xy.dat": "dec06696d2736f8c", "vel_xz.dat": "33e2cda5e1da1793",
    "vel_yz.dat": "d25ce8767bb049aa",
}


def test_solver_interface_writers_and_restarts(tmp_path):
    """hosts/solver_b200.h (the reference Solver<D>'s method names on the C ABI): every text writer byte-identical
    to the reference's, the .vts family readable, macroscopic restart (load_file) and exact restart
    (savePopulations/loadPopulations) verified inside the program."""
    out = _run("solver_selftest", [], str(tmp_path))
    assert "solver_selftest OK" in out
    for name, sha in REFERENCE_TEXT_FILES.items():
        assert _sha16(tmp_path / name) == sha, name
    a = _read_vts(tmp_path / "all.vts")
    assert list(a) == ["Velocity", "Phase", "Density", "Points"]
    assert list(_read_vts(tmp_path / "density.vts")) == ["Density", "Points"]
    assert list(_read_vts(tmp_path / "velocity.vts")) == ["Velocity", "Points"]
    assert np.array_equal(_read_vts(tmp_path / "phase.vts")["Phase"], a["Phase"])
    assert np.array_equal(_read_vts(tmp_path / "local_phase_process0.vts")["Phase"], a["Phase"])
    assert open(tmp_path / "local_phase_text_process0.dat").read() == open(tmp_path / "phase_whole.dat").read()
    # the same run through the ctypes binding
    from latticeboltzmannmethod_b200 import api
    import oracle_lib as ol
    cfg = api.microchannel_config("quarter", 14, 11, 16)
    cfg.wall_gradient = 0.02
    ctx = api.Context(cfg)
    for name, f in zip(("rho", "phase", "ux", "uy", "uz"), ol.bubble_fields(14, 11, 16, 4)):
        ctx.set_field(name, f)
    ctx.init()
    ctx.step(7)
    assert np.array_equal(ctx.get_field("phase"), a["Phase"]) and np.array_equal(ctx.get_field("rho"), a["Density"])
    assert np.array_equal(ctx.get_field("uz"), a["Velocity"].reshape(-1, 3)[:, 2])
    ctx.close()


def test_microchannel_checkpoint_restart(tmp_path):
    """LBM_CHECKPOINT / LBM_RESTART of the drop-in QuarterChannel program: 201 steps in one go equal 141 + 60 bit for bit."""
    one, two = tmp_path / "one", tmp_path / "two"
    os.mkdir(one)
    os.mkdir(two)
    _run("microchannel_quarter", [], str(one), env={"LBM_NUMTIME": "201"})
    _run("microchannel_quarter", [], str(two), env={"LBM_NUMTIME": "141", "LBM_CHECKPOINT": "ck"})
    out = _run("microchannel_quarter", [], str(two), env={"LBM_NUMTIME": "201", "LBM_RESTART": "ck"})
    assert "Time is 160" in out and "Time is 140" not in out
    a, b = _read_vts(one / "phase00200.vts"), _read_vts(two / "phase00200.vts")
    for k in ("Velocity", "Phase", "Density"):
        assert np.array_equal(a[k], b[k]), k


def test_microchannel_in_place_writes_the_same_files(tmp_path):
    """LBM_KERNEL_VARIANT=2: the drop-in QuarterChannel program on ONE population array (in-place streaming): stdout,
    every .vts and the checkpoint (taken after an odd number of steps) byte-identical to the default run; a restart
    from that checkpoint continues identically as well."""
    two, one = tmp_path / "two_arrays", tmp_path / "one_array"
    os.mkdir(two)
    os.mkdir(one)
    env = {"LBM_NUMTIME": "141", "LBM_CHECKPOINT": "ck"}
    out2 = _run("microchannel_quarter", [], str(two), env=env)
    out1 = _run("microchannel_quarter", [], str(one), env=dict(env, LBM_KERNEL_VARIANT="2"))
    assert out1 == out2
    files = sorted(os.listdir(two))
    assert files == sorted(os.listdir(one)) and "ck.pop" in files
    for f in files:
        assert open(two / f, "rb").read() == open(one / f, "rb").read(), f
    _run("microchannel_quarter", [], str(one), env={"LBM_NUMTIME": "201", "LBM_RESTART": "ck", "LBM_KERNEL_VARIANT": "2"})
    _run("microchannel_quarter", [], str(two), env={"LBM_NUMTIME": "201", "LBM_RESTART": "ck"})
    assert open(two / "phase00200.vts", "rb").read() == open(one / "phase00200.vts", "rb").read()


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("gpus", [2, 4])
def test_microchannel_on_several_gpus_writes_the_same_files(gpus, tmp_path):
    """LBM_GPUS=N: the C++ drop-in spreads its grid over N GPUs of the box (one host thread and one slab context per
    GPU, NCCL halos inside the library) -- the replacement of the reference's `mpiexec -n N` (binary.pbs:58,
    solver.hpp:24-30, :121-277).  stdout, every .vts and the checkpoint must be byte-identical to the 1-GPU run; the
    solver self-test (all writers, both restarts) must pass on N GPUs as well."""
    if _ngpu() < gpus:
        pytest.skip("needs %d GPUs" % gpus)
    one, many = tmp_path / "one", tmp_path / "many"
    os.mkdir(one)
    os.mkdir(many)
    env = {"LBM_NUMTIME": "201", "LBM_CHECKPOINT": "ck"}
    out1 = _run("microchannel_quarter", [], str(one), env=env)
    outn = _run("microchannel_quarter", [], str(many), env=dict(env, LBM_GPUS=str(gpus)))
    assert out1 == outn
    files = sorted(os.listdir(one))
    assert files == sorted(os.listdir(many)) and "phase00200.vts" in files and "ck.pop" in files
    for f in files:
        assert open(one / f, "rb").read() == open(many / f, "rb").read(), f
    # the same on ONE population array per GPU (in-place streaming, five directions per face in the halo messages)
    inplace = tmp_path / "inplace"
    os.mkdir(inplace)
    outi = _run("microchannel_quarter", [], str(inplace), env=dict(env, LBM_GPUS=str(gpus), LBM_KERNEL_VARIANT="2"))
    assert outi == out1
    for f in files:
        assert open(one / f, "rb").read() == open(inplace / f, "rb").read(), ("in place", f)
    # continue the N-GPU checkpoint on one GPU and the other way round: same files again
    for src, dst, g in ((many, one, "1"), (one, many, str(gpus))):
        _run("microchannel_quarter", [], str(dst), env={"LBM_NUMTIME": "301", "LBM_RESTART": str(src / "ck"), "LBM_GPUS": g})
    assert open(one / "phase00300.vts", "rb").read() == open(many / "phase00300.vts", "rb").read()
    st = tmp_path / "selftest"
    os.mkdir(st)
    out = _run("solver_selftest", [], str(st), env={"LBM_GPUS": str(gpus)})
    assert "solver_selftest OK" in out
    for name, sha in REFERENCE_TEXT_FILES.items():
        assert _sha16(st / name) == sha, name
