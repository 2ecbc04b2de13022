"""numpy <-> device-buffer helpers for the GPU parity tests (buffers are raw bytes,
like cl.Buffer upstream)."""
import numpy as np
import torch


def to_dev(arr, dev="cuda:0"):
    a = np.ascontiguousarray(arr)
    return torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).to(dev)


def dev_empty(nbytes, dev="cuda:0", fill=None):
    t = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=dev)
    if fill is not None:
        t.fill_(fill)
    return t


def to_host(t, dtype, shape=None):
    a = t.cpu().numpy().view(np.uint8)
    n = None if shape is None else int(np.prod(shape)) * np.dtype(dtype).itemsize
    a = a[:n] if n is not None else a[:len(a) // np.dtype(dtype).itemsize * np.dtype(dtype).itemsize]
    out = a.view(dtype)
    return out.reshape(shape) if shape is not None else out
