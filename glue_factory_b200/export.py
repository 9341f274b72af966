"""Export step after the forward path: drop-in for
`gluefactory.utils.export_predictions.export_predictions` (reference: utils/export_predictions.py:17-81).

The reference rescales the keypoints, moves every tensor of every pair to the host with its own `.cpu()` (a
device synchronisation per tensor) and writes one HDF5 group per pair; at ~1 ms per pair of forward time that loop is
the wall clock of an evaluation run.  Here a batch of predictions is rescaled, converted and packed into fixed-size
per-pair records by ONE kernel (`b200_export_pack`), copied to pinned host memory with ONE asynchronous transfer per
batch (double-buffered: batch i is written to disk while batch i+1 computes), and written as one group per pair with
the reference's dataset names (`keypoints0/1`, `keypoint_scores0/1`, `matches0/1`, `matching_scores0/1`, ...).

File format: HDF5 through h5py when it is importable (same layout as the reference: group = pair name, one dataset
per key).  h5py is absent from this image; the writer then produces an uncompressed `.npz`-style zip archive whose
members are `<pair name>/<key>.npy`, and `RecordFile` reads either form back with h5py's access pattern
(`f[name][key].__array__()`), which is what eval/hpatches.py:50-59 and models/cache_loader.py:210-264 use.
"""
from __future__ import annotations

import io
import queue
import threading
import zipfile
from pathlib import Path

import numpy as np
import torch

from . import _lib, ops

try:  # pragma: no cover - not installed in this image
    import h5py
except ImportError:
    h5py = None


def _scaled_prefix(key):
    for prefix in ("keypoints", "lines", "orig_lines"):
        if key.startswith(prefix):
            return key.replace(prefix, "")
    return None


class RecordPacker:
    """Device-side packing of a batch of predictions into [B, record_bytes] uint8 records."""

    def __init__(self, as_half=False):
        self.as_half = bool(as_half)
        self.layout = None  # [(key, np dtype, per-pair shape, offset, nbytes)]
        self._sig = None
        self.record_bytes = 0

    def _plan(self, pred):
        layout, off = [], 0
        for k, v in pred.items():
            if v.dtype == torch.float32:
                dt = np.dtype(np.float16 if self.as_half else np.float32)
            elif v.dtype == torch.int64:
                dt = np.dtype(np.int64)
            else:
                raise NotImplementedError(f"export: key {k} has dtype {v.dtype}; float32 and int64 are packed")
            shape = tuple(v.shape[1:])
            nbytes = int(np.prod(shape, dtype=np.int64)) * dt.itemsize
            off = (off + 7) // 8 * 8
            layout.append((k, dt, shape, off, nbytes))
            off += nbytes
        self.layout, self.record_bytes = layout, (off + 7) // 8 * 8
        if len(layout) > _lib.EXPORT_MAX_FIELDS:
            raise NotImplementedError(f"export: {len(layout)} keys, the packing kernel takes {_lib.EXPORT_MAX_FIELDS}")

    def pack(self, pred: dict, data: dict) -> torch.Tensor:
        """pred: {key: (B, ...) CUDA tensor}; data: the batch (for `scales`).  Returns uint8 (B, record_bytes)."""
        sig = [(k, v.dtype, tuple(v.shape[1:])) for k, v in pred.items()]
        if self.layout is None or sig != self._sig:  # keypoint counts may change from batch to batch
            self._plan(pred)
            self._sig = sig
        first = next(iter(pred.values()))
        B, dev = first.shape[0], first.device
        records = torch.empty(B, self.record_bytes, dtype=torch.uint8, device=dev)
        fields = []
        for (k, dt, shape, off, _), v in zip(self.layout, pred.values()):
            n = int(np.prod(shape, dtype=np.int64))
            idx = _scaled_prefix(k)
            if v.dtype == torch.int64:
                fields.append((v, _lib.EXPORT_I64, n, off, None))
            elif idx is not None:
                scales = data["scales"] if len(idx) == 0 else data[f"view{idx}"]["scales"]
                scales = scales.to(device=dev, dtype=torch.float32).reshape(B, 2).contiguous()
                if shape[-1] != 2:
                    raise NotImplementedError(f"export: {k} must end in an (x, y) dimension")
                fields.append((v, _lib.EXPORT_F32_KEYPOINTS, n, off, scales))
            else:
                fields.append((v, _lib.EXPORT_F32, n, off, None))
        with torch.cuda.device(dev):
            ops.export_pack(fields, B, self.record_bytes, self.as_half, records)
        return records

    def unpack(self, host_records: np.ndarray, b: int, layout=None) -> dict:
        """One pair's arrays, copied out of the host record buffer.  `layout`: the plan the records were packed
        with (defaults to the current one; the writer thread passes the snapshot taken at pack time)."""
        row = host_records[b]
        return {k: np.frombuffer(row[off: off + nbytes].tobytes(), dtype=dt).reshape(shape)
                for k, dt, shape, off, nbytes in (self.layout if layout is None else layout)}


class _Writer:
    """One group per pair.  HDF5 when h5py exists, else a zip of `<name>/<key>.npy` members."""

    def __init__(self, path):
        self.path = str(path)
        if h5py is not None:
            self.h5, self.zip = h5py.File(self.path, "w"), None
        else:
            self.h5, self.zip = None, zipfile.ZipFile(self.path, "w", zipfile.ZIP_STORED, allowZip64=True)
        self.names = set()

    def write(self, name, arrays):
        if name in self.names:  # the reference skips duplicates (RuntimeError from create_group, :76-77)
            return
        self.names.add(name)
        if self.h5 is not None:
            grp = self.h5.create_group(name)
            for k, v in arrays.items():
                grp.create_dataset(k, data=v)
            return
        for k, v in arrays.items():
            buf = io.BytesIO()
            np.save(buf, v, allow_pickle=False)
            self.zip.writestr(f"{name}/{k}.npy", buf.getvalue())

    def close(self):
        (self.h5 or self.zip).close()


class _Dataset:
    def __init__(self, arr):
        self._a = arr
        self.shape, self.dtype = arr.shape, arr.dtype

    def __array__(self, dtype=None, copy=None):
        return self._a if dtype is None else self._a.astype(dtype)

    def __getitem__(self, idx):
        return self._a[idx]


class _Group:
    def __init__(self, zf, name, keys):
        self._zf, self._name, self._keys = zf, name, keys

    def keys(self):
        return list(self._keys)

    def __contains__(self, k):
        return k in self._keys

    def __getitem__(self, k):
        if k not in self._keys:
            raise KeyError(k)
        with self._zf.open(f"{self._name}/{k}.npy") as f:
            return _Dataset(np.load(io.BytesIO(f.read()), allow_pickle=False))


class RecordFile:
    """Read access with h5py's pattern: `RecordFile(path)[pair_name][key][...]`."""

    def __init__(self, path, mode="r"):
        assert mode == "r"
        self._h5 = h5py.File(str(path), "r") if h5py is not None and not zipfile.is_zipfile(str(path)) else None
        if self._h5 is None:
            self._zf = zipfile.ZipFile(str(path), "r")
            self._groups = {}
            for member in self._zf.namelist():
                name, key = member[:-4].rsplit("/", 1)
                self._groups.setdefault(name, []).append(key)

    def keys(self):
        return list(self._h5.keys()) if self._h5 is not None else list(self._groups)

    def __contains__(self, name):
        return name in (self._h5 if self._h5 is not None else self._groups)

    def __getitem__(self, name):
        if self._h5 is not None:
            return self._h5[name]
        return _Group(self._zf, name, self._groups[name])

    def close(self):
        (self._h5 or self._zf).close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


@torch.no_grad()
def export_predictions(loader, model, output_file, as_half=False, keys="*", callback_fn=None, optional_keys=[]):
    """Same contract as the reference; batches of any size; one D2H copy per batch, written by a host thread."""
    from gluefactory.utils.tensor import batch_to_device

    assert keys == "*" or isinstance(keys, (tuple, list))
    if not torch.cuda.is_available():
        raise RuntimeError("glue_factory_b200 export_predictions runs on CUDA only (no CPU fallback)")
    Path(output_file).parent.mkdir(exist_ok=True, parents=True)
    writer = _Writer(output_file)
    device = torch.device("cuda", torch.cuda.current_device())
    model = model.to(device).eval()
    packer = RecordPacker(as_half)
    slots = [dict(host=None, done=torch.cuda.Event(), busy=threading.Event()) for _ in range(2)]
    for s in slots:
        s["busy"].set()  # free
    jobs: queue.Queue = queue.Queue()
    errors = []

    def drain():
        while True:
            job = jobs.get()
            if job is None:
                return
            slot, names, layout = job
            try:
                slot["done"].synchronize()  # this batch's copy has landed; later batches keep running
                host = slot["host"].numpy()
                for b, name in enumerate(names):
                    writer.write(name, packer.unpack(host, b, layout))
            except Exception as e:  # surfaced to the caller after the loop
                errors.append(e)
            finally:
                slot["busy"].set()

    thread = threading.Thread(target=drain, daemon=True)
    thread.start()
    i = 0
    for data_ in loader:
        data = batch_to_device(data_, device, non_blocking=True)
        pred = model(data)
        if callback_fn is not None:
            pred = {**callback_fn(pred, data), **pred}
        if keys != "*":
            if len(set(keys) - set(pred.keys())) > 0:
                raise ValueError(f"Missing key {set(keys) - set(pred.keys())}")
            pred = {k: v for k, v in pred.items() if k in list(keys) + list(optional_keys)}
        assert len(pred) > 0
        records = packer.pack(pred, data)
        slot = slots[i & 1]
        slot["busy"].wait()   # the writer has finished with this pinned buffer
        slot["busy"].clear()
        if slot["host"] is None or slot["host"].shape != records.shape:
            slot["host"] = torch.empty(records.shape, dtype=torch.uint8).pin_memory()
        slot["host"].copy_(records, non_blocking=True)
        slot["done"].record()
        names = data["name"] if isinstance(data["name"], (list, tuple)) else [data["name"]]
        jobs.put((slot, list(names), list(packer.layout)))
        i += 1
        del pred
    jobs.put(None)
    thread.join()
    writer.close()
    if errors:
        raise errors[0]
    return output_file
