"""Lazy checkpoint streaming with the one-time weight repack inside the load.

ref: /root/reference/lit_llama/utils.py:198-376 (``NotYetLoadedTensor``, ``LazyLoadingUnpickler``, ``lazy_load``): the
reference unpickles a ``torch.save`` archive into placeholders and lets ``torch._C.PyTorchFileReader`` materialise one
tensor when torch first touches it, so that ``model.load_state_dict`` never holds the whole checkpoint on the host.

Same surface here (``with lazy_load(path) as sd: model.load_state_dict(sd)``; placeholders answer ``shape`` / ``dtype`` /
``size()`` / ``contiguous()`` and materialise under any torch function), built for a B200 box instead of translated:

* the archive is INDEXED, not read: ``torch.save`` stores every storage as an uncompressed, 64-byte-aligned zip record, so
  a tensor is (file offset, dtype, shape, strides) and its bytes can be read with one ``os.preadv`` - no zip layer, no
  intermediate ``bytes`` object;
* ``stream_load_state_dict`` moves a checkpoint disk -> pinned staging buffer -> HBM on a copy stream with two staging
  buffers in flight (the read of tensor i+1 overlaps the H2D copy of tensor i), and runs each quantized Linear's ONE-TIME
  repack (``ColBlockQuantizedLinear.prepare`` -> ``ll_int4_prepare``: checkpoint bytes -> the GEMV's fragment-ordered
  stream) on the compute stream as soon as that layer's three tensors have landed, i.e. under the next layer's disk read.
  Peak host memory is the two staging buffers, whatever the checkpoint size (65B int4: 34.5 GB through 2 x 256 MB).

The unpickler resolves only the handful of globals a state dict needs (tensor rebuild functions, storage classes,
``OrderedDict``); anything else raises.
"""
from __future__ import annotations

import io
import os
import pickle
import struct
import zipfile
from collections import OrderedDict
from typing import Dict, Optional

import numpy as np
import torch

_STORAGE_DTYPES = {
    "FloatStorage": torch.float32, "DoubleStorage": torch.float64, "HalfStorage": torch.float16,
    "BFloat16Storage": torch.bfloat16, "LongStorage": torch.int64, "IntStorage": torch.int32,
    "ShortStorage": torch.int16, "CharStorage": torch.int8, "ByteStorage": torch.uint8, "BoolStorage": torch.bool,
}
# numpy has no bfloat16: bytes are read as the same-sized integer type and re-viewed
_NP_CARRIER = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}


class _StorageRef:
    """One storage record of the archive: where its bytes are, not the bytes."""

    __slots__ = ("key", "dtype", "numel", "offset")

    def __init__(self, key: str, dtype: torch.dtype, numel: int, offset: int):
        self.key, self.dtype, self.numel, self.offset = key, dtype, numel, offset


class _StorageClass:
    def __init__(self, dtype):
        self.dtype = dtype


class _Archive:
    """Index of a ``torch.save`` zip archive: record name -> absolute offset of its (stored) bytes."""

    def __init__(self, path):
        self.path = str(path)
        self.fd = os.open(self.path, os.O_RDONLY)
        self.offsets: Dict[str, int] = {}
        self.sizes: Dict[str, int] = {}
        with zipfile.ZipFile(self.path) as zf:
            self.prefix = None
            for info in zf.infolist():
                if info.compress_type != zipfile.ZIP_STORED:
                    raise ValueError(f"{self.path}: record {info.filename} is compressed; torch.save never does that")
                # local file header: 30 fixed bytes, then name and extra field (their lengths at bytes 26..29)
                hdr = os.pread(self.fd, 30, info.header_offset)
                if hdr[:4] != b"PK\x03\x04":
                    raise ValueError(f"{self.path}: bad local header for {info.filename}")
                nlen, xlen = struct.unpack("<HH", hdr[26:30])
                self.offsets[info.filename] = info.header_offset + 30 + nlen + xlen
                self.sizes[info.filename] = info.file_size
                if info.filename.endswith("/data.pkl"):
                    self.prefix = info.filename[: -len("data.pkl")]
            if self.prefix is None:
                raise ValueError(f"{self.path}: not a torch.save zip archive (no data.pkl)")
            self.pickle_bytes = zf.read(self.prefix + "data.pkl")

    def record(self, key: str):
        name = f"{self.prefix}data/{key}"
        return self.offsets[name], self.sizes[name]

    def read_into(self, buf: memoryview, offset: int) -> None:
        """Fill ``buf`` from absolute file offset ``offset`` (straight into the caller's memory)."""
        done, n = 0, len(buf)
        while done < n:
            got = os.preadv(self.fd, [buf[done:]], offset + done)
            if got <= 0:
                raise IOError(f"{self.path}: short read at {offset + done}")
            done += got

    def close(self):
        if self.fd is not None:
            os.close(self.fd)
            self.fd = None


class NotYetLoadedTensor:
    """Placeholder for one tensor of the archive (ref utils.py:198-336).  Materialises on ``_load_tensor()``, on
    ``contiguous()`` / ``to()`` / ``float()`` ..., and whenever it is an argument of a torch function (``param.copy_(x)``
    inside ``load_state_dict``)."""

    is_meta = False

    def __init__(self, archive: _Archive, storage: _StorageRef, storage_offset: int, size, stride, requires_grad=False,
                 as_parameter=False):
        self._archive, self._storage = archive, storage
        self._offset, self._size, self._stride = int(storage_offset), tuple(int(s) for s in size), tuple(int(s) for s in stride)
        self.requires_grad = bool(requires_grad)
        self._as_parameter = as_parameter

    # ---- metadata, no I/O
    @property
    def shape(self):
        return torch.Size(self._size)

    @property
    def dtype(self):
        return self._storage.dtype

    @property
    def ndim(self):
        return len(self._size)

    @property
    def device(self):
        return torch.device("cpu")

    def size(self, dim: Optional[int] = None):
        return self.shape if dim is None else self._size[dim]

    def dim(self):
        return len(self._size)

    def numel(self):
        n = 1
        for s in self._size:
            n *= s
        return n

    def stride(self, dim: Optional[int] = None):
        return self._stride if dim is None else self._stride[dim]

    def span(self):
        """(first element, element count) of the storage range this view touches."""
        if self.numel() == 0:
            return self._offset, 0
        return self._offset, 1 + sum((s - 1) * st for s, st in zip(self._size, self._stride))

    def nbytes_span(self):
        return self.span()[1] * self.dtype.itemsize

    def file_offset(self):
        base, _ = self._archive.record(self._storage.key)
        return base + self._offset * self.dtype.itemsize

    # ---- materialisation
    def read_span_into(self, buf: memoryview) -> None:
        """The bytes of ``span()`` into ``buf`` (``nbytes_span()`` long)."""
        if len(buf):
            self._archive.read_into(buf, self.file_offset())

    def view_of_span(self, flat: torch.Tensor) -> torch.Tensor:
        """This tensor as a strided view of ``flat`` (1-D, this dtype, holding ``span()``)."""
        return torch.as_strided(flat, self._size, self._stride, 0)

    def _load_tensor(self) -> torch.Tensor:
        isz = self.dtype.itemsize
        n = self.span()[1]
        carrier = np.empty(n, dtype=_NP_CARRIER[isz])
        self.read_span_into(memoryview(carrier).cast("B"))
        flat = torch.from_numpy(carrier).view(self.dtype) if n else torch.empty(0, dtype=self.dtype)
        t = self.view_of_span(flat)
        if self._as_parameter:
            t = torch.nn.Parameter(t, requires_grad=self.requires_grad)
        return t

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        args = [a._load_tensor() if isinstance(a, NotYetLoadedTensor) else a for a in args]
        kwargs = {k: (v._load_tensor() if isinstance(v, NotYetLoadedTensor) else v) for k, v in kwargs.items()}
        return func(*args, **kwargs)

    def __getattr__(self, name):
        # anything else a caller asks for (contiguous, to, float, t, ...) is answered by the materialised tensor
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self._load_tensor(), name)

    def __repr__(self):
        return f"NotYetLoadedTensor(shape={tuple(self._size)}, dtype={self.dtype}, record={self._storage.key})"


class LazyLoadingUnpickler(pickle.Unpickler):
    """ref utils.py:339-361: tensor rebuild calls become placeholders; storages become references into the archive."""

    def __init__(self, file, archive: _Archive):
        super().__init__(file)
        self.archive = archive

    def find_class(self, module, name):
        if module == "torch._utils" and name == "_rebuild_tensor_v2":
            return self._rebuild_tensor_v2
        if module == "torch._utils" and name == "_rebuild_parameter":
            return self._rebuild_parameter
        if module == "torch._tensor" and name == "_rebuild_from_type_v2":
            return self._rebuild_from_type_v2
        if module == "collections" and name == "OrderedDict":
            return OrderedDict
        if module == "torch" and name in _STORAGE_DTYPES:
            return _StorageClass(_STORAGE_DTYPES[name])
        if module == "torch" and name in ("Size",):
            return torch.Size
        if module == "torch" and name in ("float32", "float16", "bfloat16", "int8", "uint8", "int32", "int64", "bool", "float64"):
            return getattr(torch, name)
        raise pickle.UnpicklingError(f"lazy_load: global {module}.{name} is not part of a state dict")

    def persistent_load(self, pid):
        kind, cls, key, _location, numel = pid
        if kind != "storage" or not isinstance(cls, _StorageClass):
            raise pickle.UnpicklingError(f"lazy_load: unexpected persistent id {pid!r}")
        return _StorageRef(str(key), cls.dtype, int(numel), 0)

    def _rebuild_tensor_v2(self, storage, storage_offset, size, stride, requires_grad=False, backward_hooks=None, metadata=None):
        return NotYetLoadedTensor(self.archive, storage, storage_offset, size, stride, requires_grad)

    @staticmethod
    def _rebuild_parameter(data, requires_grad, backward_hooks):
        if isinstance(data, NotYetLoadedTensor):
            data.requires_grad, data._as_parameter = bool(requires_grad), True
            return data
        return torch._utils._rebuild_parameter(data, requires_grad, backward_hooks)

    @staticmethod
    def _rebuild_from_type_v2(func, new_type, args, state):
        ret = func(*args)
        if isinstance(ret, NotYetLoadedTensor):
            return ret                      # tensor subclasses carry no data of their own in a state dict
        return torch._tensor._rebuild_from_type_v2(lambda: ret, new_type, (), state)


class lazy_load:
    """ref utils.py:364-376: ``with lazy_load(path) as state_dict`` - values are ``NotYetLoadedTensor``."""

    def __init__(self, fn):
        self.archive = _Archive(fn)
        self.sd = LazyLoadingUnpickler(io.BytesIO(self.archive.pickle_bytes), self.archive).load()

    def __enter__(self):
        return self.sd

    def __exit__(self, exc_type, exc_val, exc_tb):
        self.archive.close()
        self.sd = None


# ----------------------------------------------------------------------------- disk -> pinned -> HBM, repack under the read
class _Stager:
    """Two pinned staging buffers and a copy stream: ``fetch`` returns a device tensor whose H2D copy is in flight on the
    copy stream (the caller's stream is made to wait for it); the read of the next tensor overlaps that copy."""

    def __init__(self, device, buffer_bytes: int):
        self.device = torch.device(device)
        self.nbuf = 2
        self.bytes = int(buffer_bytes)
        self.pinned = [torch.empty(self.bytes, dtype=torch.uint8).pin_memory() for _ in range(self.nbuf)]
        self.events = [None] * self.nbuf
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.turn = 0
        self.h2d_bytes = 0

    def fetch(self, lt: NotYetLoadedTensor) -> torch.Tensor:
        nbytes = lt.nbytes_span()
        if nbytes > self.bytes:          # a tensor larger than a staging buffer: grow both (rare: wte / lm_head in fp32)
            torch.cuda.synchronize(self.device)
            self.bytes = nbytes
            self.pinned = [torch.empty(self.bytes, dtype=torch.uint8).pin_memory() for _ in range(self.nbuf)]
            self.events = [None] * self.nbuf
        i = self.turn
        self.turn = (i + 1) % self.nbuf
        if self.events[i] is not None:
            self.events[i].synchronize()       # the copy that last used this buffer has left it
        host = self.pinned[i][:nbytes]
        lt.read_span_into(memoryview(host.numpy()))
        with torch.cuda.stream(self.copy_stream):
            dev = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            dev.copy_(host, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        self.events[i] = ev
        self.h2d_bytes += nbytes
        torch.cuda.current_stream(self.device).wait_event(ev)
        dev.record_stream(torch.cuda.current_stream(self.device))
        return lt.view_of_span(dev.view(lt.dtype))


def stream_load_state_dict(model: torch.nn.Module, path, strict: bool = True, prepare_batch: Optional[int] = 1,
                           buffer_bytes: int = 256 << 20) -> dict:
    """Load ``path`` into ``model`` (already on its CUDA device, e.g. built under ``EmptyInitOnDevice`` +
    ``quantization(...)``) tensor by tensor through the staging pipeline, and build every quantized Linear's GEMV stream
    for batch ``prepare_batch`` as soon as its tensors are in (``None``: no repack).  Returns counters."""
    from .quantization import ColBlockQuantizedLinear

    params = dict(model.named_parameters())
    params.update(dict(model.named_buffers()))
    device = next(iter(params.values())).device
    if device.type != "cuda":
        raise RuntimeError("stream_load_state_dict needs the model on a CUDA device: there is no CPU path")
    qlins = {name: m for name, m in model.named_modules() if isinstance(m, ColBlockQuantizedLinear)}
    pending = {name: {f"{name}.{k}" if name else k for k in ("quant_weight", "scales", "zeros")} for name in qlins}
    stager = _Stager(device, buffer_bytes)
    missing, unexpected, repacked = [], [], 0
    with lazy_load(path) as sd, torch.no_grad():
        for name, m in qlins.items():      # grouped checkpoints re-shape a per-row module before its tensors arrive
            s = sd.get(f"{name}.scales" if name else "scales")
            if s is not None and tuple(s.shape) != tuple(m.scales.shape) and s.shape[0] == m.out_features and s.shape[1] > 0 \
                    and m.in_features % s.shape[1] == 0:
                m.tile_cols = m.in_features // s.shape[1]
                m.scales = torch.empty((m.out_features, s.shape[1]), dtype=m.scales.dtype, device=device)
                m.zeros = torch.empty_like(m.scales)
                params[f"{name}.scales"], params[f"{name}.zeros"] = m.scales, m.zeros
        for key, lt in sd.items():
            dst = params.get(key)
            if dst is None:
                unexpected.append(key)
                continue
            if tuple(dst.shape) != tuple(lt.shape):
                raise RuntimeError(f"size mismatch for {key}: checkpoint {tuple(lt.shape)}, model {tuple(dst.shape)}")
            dst.copy_(stager.fetch(lt))
            owner = key.rsplit(".", 1)[0] if "." in key else ""
            if owner in pending:
                pending[owner].discard(key)
                if not pending[owner]:
                    lin = qlins[owner]
                    lin._streams = {}
                    if prepare_batch is not None and lin.prepare(prepare_batch) is not None:
                        repacked += 1
        missing = [k for k in params if k not in sd]
    if strict and (missing or unexpected):
        raise RuntimeError(f"stream_load_state_dict: missing keys {missing[:5]}, unexpected keys {unexpected[:5]}")
    torch.cuda.current_stream(device).synchronize()
    return {"h2d_bytes": stager.h2d_bytes, "repacked_layers": repacked, "missing": missing, "unexpected": unexpected,
            "staging_bytes": stager.nbuf * stager.bytes}
