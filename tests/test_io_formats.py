"""On-disk formats: the drop-in's writers produce, byte for byte, the files the
reference's own writers produced for the same arrays (tests/golden/io/, written by
oracle/gen_golden_io.py from /root/reference/src/LBM/utils/io.py), for numpy and torch inputs; the
readers invert them; record / balance lines equal the demo's formatting; the field cache round-trips."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

IO = os.path.join(GOLDEN, "io")


def _inputs(kind):
    with np.load(os.path.join(IO, "inputs.npz")) as z:
        a = {k: z[k] for k in z.files}
    return a if kind == "np" else {k: torch.from_numpy(v) for k, v in a.items()}


def _same(path_a, path_b):
    with open(path_a, "rb") as fa, open(path_b, "rb") as fb:
        a, b = fa.read(), fb.read()
    assert a == b, "{} differs from {}".format(path_a, path_b)


@pytest.mark.parametrize("kind", ["np", "torch"])
def test_plt_writers_match_reference_bytes(tmp_path, kind):
    from src.LBM.utils import io
    d = _inputs(kind)
    out = str(tmp_path / "o.dat")
    io.dump_2d_plt_file_C_rho(out, d["C"], d["rho"], d["u"], d["v"], 0, 0)
    _same(out, os.path.join(IO, "c_rho_staggered_%s.dat" % kind))
    io.dump_2d_plt_file_C_rho(out, d["C"], d["rho"], d["uc"], d["vc"], 0, 0)
    _same(out, os.path.join(IO, "c_rho_centred_%s.dat" % kind))
    io.dump_2d_plt_file_single(out, d["rho"], d["u"], d["v"], 0, 0)
    _same(out, os.path.join(IO, "single_%s.dat" % kind))
    io.dump_2d_plt_file_balance(out, d["C"], d["vel3"], d["h"], 0)
    _same(out, os.path.join(IO, "balance_%s.dat" % kind))


def test_plt_readers_match_reference():
    from src.LBM.utils import io
    a = _inputs("np")
    with np.load(os.path.join(IO, "read_back.npz")) as z:
        ref = {k: z[k] for k in z.files}
    C0, vel0, h0 = np.zeros_like(a["C"]), np.zeros((1, 2) + a["C"].shape[2:], np.float32), np.zeros_like(a["h"])
    C1, vel1, h1 = io.read_2d_plt_file_balance(os.path.join(IO, "balance_np.dat"), C0, vel0, h0, 0)
    assert np.array_equal(C1, ref["bal_C"]) and np.array_equal(vel1, ref["bal_vel"]) and np.array_equal(h1, ref["bal_h"])
    assert np.array_equal(h1, a["h"])  # text round trip of float32 is exact (repr of the double is exact)
    C, rho, u, v = io.read_2d_plt_file_C_rho(os.path.join(IO, "c_rho_centred_np.dat"))
    for got, key in ((C, "C"), (rho, "rho"), (u, "u"), (v, "v")):
        assert got.dtype == np.float32 and np.array_equal(got, ref[key])
    with pytest.raises(RuntimeError, match="does not match"):
        io.read_2d_plt_file_balance(os.path.join(IO, "balance_np.dat"), np.zeros((1, 1, 3, 3), np.float32), vel0, h0, 0)


@pytest.mark.parametrize("as_torch", [False, True])
def test_record_and_balance_lines(tmp_path, as_torch):
    import jrl_io
    with np.load(os.path.join(IO, "record_inputs.npz")) as z:
        rec, dtc, vor = z["rec"], float(z["dt_conv_real"]), float(z["vel_obs_real"])
    r = torch.from_numpy(rec) if as_torch else rec
    n = rec.shape[0]
    out = str(tmp_path / "record.txt")
    jrl_io.write_record_lines(out, [10 * i + 9 for i in range(3)], r[:3], dtc, vor, header=True)
    jrl_io.write_record_lines(out, [10 * i + 9 for i in range(3, n)], r[3:], dtc, vor)      # appended, as the demo does
    _same(out, os.path.join(IO, "record.txt"))
    out = str(tmp_path / "balance.txt")
    jrl_io.write_balance_lines(out, [10 * i for i in range(n)], r[:, 0], dtc, header=True)
    _same(out, os.path.join(IO, "balance.txt"))
    with pytest.raises(RuntimeError, match="steps"):
        jrl_io.write_record_lines(out, [1, 2], r[:3], dtc, vor)


def test_field_cache_round_trip(tmp_path):
    import jrl_io
    g = torch.Generator().manual_seed(2)
    C, h = torch.rand((1, 1, 6, 4), generator=g), torch.rand((1, 9, 6, 4), generator=g)
    name = str(tmp_path / "sub" / "C_balance.h5")
    assert not jrl_io.fields_exist(name)
    with pytest.raises(FileNotFoundError):
        jrl_io.load_fields(name)
    written = jrl_io.save_fields(name, C=C, h=h.numpy())
    assert os.path.exists(written) and jrl_io.fields_exist(name)
    assert written == (name if jrl_io.HAVE_H5PY else name + ".npz")
    back = jrl_io.load_fields(name, names=["C", "h"], dtype=torch.float32)
    assert torch.equal(back["C"], C) and torch.equal(back["h"], h)
    assert set(jrl_io.load_fields(name)) == {"C", "h"}
    with pytest.raises(KeyError):
        jrl_io.load_fields(name, names=["vel"])


def test_snapshotter_needs_cuda():
    import jrl_io
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError, match="no CPU path"):
        jrl_io.AsyncSnapshotter()


@pytest.mark.gpu
def test_async_snapshots_equal_synchronous_files(tmp_path):
    """Snapshots streamed through the pinned ring while the producing stream keeps overwriting the fields hold
    exactly the values of the step they were submitted at."""
    import jrl_io
    from src.LBM.utils import io
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(4)
    base = {k: torch.rand(s, generator=g).to(dev) for k, s in
            (("C", (1, 1, 40, 24)), ("phi", (1, 1, 40, 24)), ("u", (1, 1, 40, 25)), ("v", (1, 1, 41, 24)))}
    snap = jrl_io.AsyncSnapshotter(depth=2)
    want = []
    for step in range(5):
        for t in base.values():
            t.mul_(1.0009765625).add_(0.03125 * step)        # in-place update on the producing stream
        path = str(tmp_path / ("%03d.dat" % step))
        want.append({k: t.clone() for k, t in base.items()})
        snap.submit(lambda h, p=path: io.dump_2d_plt_file_C_rho(p, h["C"], h["phi"], h["u"], h["v"], 0, 0), **base)
    snap.close()
    assert snap.snapshots == 5 and snap.bytes_copied == 5 * sum(t.numel() * 4 for t in base.values())
    for step, w in enumerate(want):
        ref = str(tmp_path / ("ref_%03d.dat" % step))
        io.dump_2d_plt_file_C_rho(ref, w["C"], w["phi"], w["u"], w["v"], 0, 0)
        _same(str(tmp_path / ("%03d.dat" % step)), ref)


@pytest.mark.gpu
def test_snapshot_writer_errors_surface(tmp_path):
    import jrl_io
    snap = jrl_io.AsyncSnapshotter(depth=1)

    def boom(h):
        raise ValueError("disk full")
    snap.submit(boom, x=torch.zeros(4, device="cuda:0"))
    with pytest.raises(RuntimeError, match="snapshot writer failed"):
        snap.flush()
    snap.close()
