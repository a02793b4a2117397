"""On-disk formats either side of the hot path: the field caches the demos
pass between stages, their `record_*.txt` / `balance.txt` series, and snapshot streaming off the
device while the step loop keeps running.

  * `save_fields` / `load_fields`   -- the `vel` / `C` / `h` caches (demo_2d_LBM_JRL_fluid_cylinder.py:266-268,
    demo_2d_LBM_JRL_C_cylinder.py:209-213, 251-254, 276).  The reference writes HDF5 through h5py; h5py is used
    here when it is importable, else the same dataset names go into an `.npz` next to the requested
    path (`<name>.h5.npz`).  `load_fields` reads whichever exists.
  * `write_record_lines` / `write_balance_lines` -- the text series, byte for byte as the demo formats
    them (:245-250, :270-272, :383-393), from the `records` tensor `jrl_conc_walk_f32` fills on the device:
    one device->host copy for the whole walk instead of four `.item()` synchronisations per sample.
  * `AsyncSnapshotter` -- copies fields into a ring of pinned host buffers on its own CUDA stream (ordered after
    the producing stream by an event, so the step loop never waits for PCIe), and hands them to a writer thread
    that produces the `.dat` / `.png` / cache files.  `flush()` joins everything.

Nothing here is on the per-step path and nothing here computes lattice quantities: inputs are tensors the
library produced."""
import os
import queue
import threading

import numpy as np
import torch

try:  # optional: absent from the build image
    import h5py  # noqa: F401
    HAVE_H5PY = True
except Exception:  # pragma: no cover - depends on the image
    h5py = None
    HAVE_H5PY = False


# ---------------------------------------------------------------------------------------------
# field caches
# ---------------------------------------------------------------------------------------------
def _npz_path(filename):
    return filename + ".npz"


def _is_npz(filename):
    try:
        with open(filename, "rb") as f:
            return f.read(2) == b"PK"
    except OSError:
        return False


def fields_exist(filename):
    """True if a cache written by `save_fields` (either container) is present: the demos' `os.path.exists` test."""
    return (HAVE_H5PY and os.path.exists(filename)) or os.path.exists(_npz_path(filename))


def save_fields(filename, **fields):
    """Write named fields (torch tensors or numpy arrays) as datasets of one cache file.  Returns the path written."""
    host = {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in fields.items()}
    os.makedirs(os.path.dirname(os.path.abspath(filename)), exist_ok=True)
    if HAVE_H5PY:
        with h5py.File(filename, "w") as f:
            for k, v in host.items():
                f.create_dataset(k, data=v)
        return filename
    tmp = _npz_path(filename) + ".tmp.npz"
    np.savez(tmp, **host)
    os.replace(tmp, _npz_path(filename))  # a reader never sees a half-written cache
    return _npz_path(filename)


def load_fields(filename, names=None, device=None, dtype=None):
    """Read datasets back as torch tensors (`torch.from_numpy(np.asarray(h5_data[name])).to(dtype).to(device)` in the
    demos).  Raises FileNotFoundError when neither container exists, KeyError for a missing dataset."""
    if HAVE_H5PY and os.path.exists(filename):
        with h5py.File(filename, "r") as f:
            data = {k: np.asarray(f[k]) for k in (names or f.keys())}
    elif os.path.exists(_npz_path(filename)) or _is_npz(filename):
        # `<name>.h5.npz` (save_fields without h5py), or an .npz container AT the requested path: what the demos
        # themselves write through the h5py stand-in used where h5py is not installed (tests/standins/h5py.py)
        with np.load(filename if _is_npz(filename) else _npz_path(filename)) as z:
            data = {k: z[k] for k in (names or z.files)}
    else:
        raise FileNotFoundError("no field cache at {} (or {})".format(filename, _npz_path(filename)))
    out = {}
    for k, v in data.items():
        t = torch.from_numpy(np.ascontiguousarray(v))
        if dtype is not None:
            t = t.to(dtype)
        out[k] = t.to(device) if device is not None else t
    return out


# ---------------------------------------------------------------------------------------------
# record / balance series
# ---------------------------------------------------------------------------------------------
RECORD_HEADER = "Step Inflow Current u, v\n"      # demo_2d_LBM_JRL_C_cylinder.py:272
BALANCE_HEADER = "Step Inflow\n"                  # :189


def _rows(records):
    r = records.detach().cpu().numpy() if isinstance(records, torch.Tensor) else np.asarray(records)
    return np.asarray(r, dtype=np.float64).tolist()   # float32 -> the Python floats `.item()` returns


def write_record_lines(record_name, steps, records, dt_conv_real, vel_obs_real, header=False):
    """Append one `record_*.txt` line per sample: `(step + 1) * dt_conv_real * vel_obs_real * 1e6` and the four
    recorded sums (inflow, current, u, v), formatted as the demo does (:383-393).  `steps[i]` is the loop index of
    sample i, `records[i]` its row of the tensor `conc_walk` returns."""
    rows = _rows(records)
    if len(rows) != len(steps):
        raise RuntimeError("write_record_lines: {} steps for {} record rows".format(len(steps), len(rows)))
    with open(record_name, "w" if header else "a+") as fo:
        if header:
            fo.write(RECORD_HEADER)
        for s, r in zip(steps, rows):
            fo.write("{} {} {} {} {}\n".format((int(s) + 1) * dt_conv_real * vel_obs_real * 1e6, r[0], r[1], r[2], r[3]))


def write_balance_lines(balance_name, steps, inflow, dt_conv_real, header=False):
    """Append `balance.txt` lines `(step + 1) * dt_conv_real  inflow` (:245-250)."""
    vals = _rows(inflow)
    vals = [v[0] if isinstance(v, list) else v for v in vals]
    if len(vals) != len(steps):
        raise RuntimeError("write_balance_lines: {} steps for {} values".format(len(steps), len(vals)))
    with open(balance_name, "w" if header else "a+") as fo:
        if header:
            fo.write(BALANCE_HEADER)
        for s, v in zip(steps, vals):
            fo.write("{} {}\n".format((int(s) + 1) * dt_conv_real, v))


# ---------------------------------------------------------------------------------------------
# snapshot streaming
# ---------------------------------------------------------------------------------------------
class AsyncSnapshotter:
    """Stream snapshots off the device without stalling the step loop.

        snap = AsyncSnapshotter(depth=2)
        for step in ...:
            ... launch the step ...
            if step % interval == 0:
                snap.submit(lambda h: dump_2d_plt_file_C_rho(path, h["C"], h["phi"], h["u"], h["v"], 0, 0),
                            C=C, phi=phi_obs, u=u, v=v)
        snap.flush()

    `submit` copies the fields device-to-device into per-slot staging buffers on the CURRENT stream (in order with the
    step loop, HBM speed), records an event, makes the copy stream wait for it, enqueues the staging->pinned copies there
    and returns; the writer thread synchronises on the copy's own event and then calls `fn` with a dict of numpy views
    of the pinned buffers.  The producing stream never waits for PCIe or for the disk, and the source tensors may be
    overwritten as soon as `submit` has returned (the step loop's ping-pong buffers are, two steps later).  `depth`
    snapshots can be in flight; `submit` blocks the host (not the GPU) when all slots are still being written out."""

    def __init__(self, depth=2, device=None, guard=True):
        if not torch.cuda.is_available():
            raise RuntimeError("AsyncSnapshotter needs a CUDA device (no CPU path)")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.depth = int(depth)
        self.guard = guard  # kept for call compatibility; unused since snapshots are staged on the device
        self.stream = torch.cuda.Stream(device=self.device)
        self._slots = [dict(bufs={}, dev={}, free=threading.Event()) for _ in range(self.depth)]
        for s in self._slots:
            s["free"].set()
        self._next = 0
        self._q = queue.Queue()
        self._err = None
        self.bytes_copied = 0
        self.snapshots = 0
        self._thread = threading.Thread(target=self._work, name="jrl-snapshot-writer", daemon=True)
        self._thread.start()

    def _pinned(self, slot, name, t):
        b = slot["bufs"].get(name)
        if b is None or b.shape != t.shape or b.dtype != t.dtype:
            b = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            slot["bufs"][name] = b
        return b

    def submit(self, fn, **tensors):
        if self._err is not None:
            raise RuntimeError("snapshot writer failed") from self._err
        slot = self._slots[self._next]
        self._next = (self._next + 1) % self.depth
        slot["free"].wait()
        slot["free"].clear()
        # Snapshot = a device-to-device copy into this slot's staging buffers ON THE PRODUCING STREAM (HBM speed, in
        # order with the step loop, so the loop may overwrite its ping-pong buffers right after), then the D2H copy of
        # the staging buffers on the copy stream.  The producing stream never waits for PCIe or the disk; a slot is
        # reused only after the writer thread has finished with it (slot["free"]).
        cur = torch.cuda.current_stream(self.device)
        staged = {}
        for name, t in tensors.items():
            t = t.detach()
            d = slot["dev"].get(name)
            if d is None or d.shape != t.shape or d.dtype != t.dtype:
                d = slot["dev"][name] = torch.empty(t.shape, dtype=t.dtype, device=self.device)
            d.copy_(t, non_blocking=True)
            staged[name] = d
        ready = torch.cuda.Event()
        ready.record(cur)
        self.stream.wait_event(ready)
        host = {}
        with torch.cuda.stream(self.stream):
            for name, d in staged.items():
                b = self._pinned(slot, name, d)
                b.copy_(d, non_blocking=True)
                host[name] = b
                self.bytes_copied += d.numel() * d.element_size()
            done = torch.cuda.Event()
            done.record(self.stream)
        self.snapshots += 1
        self._q.put((fn, host, done, slot))

    def _work(self):
        while True:
            item = self._q.get()
            if item is None:
                return
            fn, host, done, slot = item
            try:
                done.synchronize()
                fn({k: v.numpy() for k, v in host.items()})
            except BaseException as e:  # surfaced by the next submit / flush
                self._err = e
            finally:
                slot["free"].set()
                self._q.task_done()

    def flush(self):
        self._q.join()
        if self._err is not None:
            err, self._err = self._err, None
            raise RuntimeError("snapshot writer failed") from err

    def close(self):
        self.flush()
        self._q.put(None)
        self._thread.join()
