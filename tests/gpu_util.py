"""Helpers for the -m gpu tests: column-major device buffers through torch (plumbing only)."""
import numpy as np
import torch

NP = {"s": np.float32, "d": np.float64, "c": np.complex64, "z": np.complex128}
TORCH = {"s": torch.float32, "d": torch.float64, "c": torch.complex64, "z": torch.complex128}


class DevMat:
    """Column-major m x n matrix with leading dimension ld on cuda:0, backed by a flat torch tensor."""

    def __init__(self, a, ld=None, slack=32):
        a = np.asarray(a)
        self.m, self.n = a.shape
        self.ld = max(1, self.m) if ld is None else ld
        self.dtype = a.dtype
        host = np.full((max(1, self.n), self.ld), -7.7, dtype=a.dtype)  # host[j, i] = a[i, j]
        if a.size:
            host[: self.n, : self.m] = a.T
        flat = np.concatenate([host.reshape(-1), np.full(slack, -7.7, dtype=a.dtype)])
        self.t = torch.from_numpy(flat).cuda()

    @property
    def ptr(self):
        return self.t.data_ptr()

    def at(self, i, j):
        return self.ptr + (i + j * self.ld) * self.t.element_size()

    def get(self):
        h = self.t.cpu().numpy()[: self.ld * max(1, self.n)].reshape(max(1, self.n), self.ld)
        return np.ascontiguousarray(h[: self.n, : self.m].T)

    def pad(self):
        """The padding rows (ld > m) as an array: must stay at the fill value."""
        h = self.t.cpu().numpy()[: self.ld * max(1, self.n)].reshape(max(1, self.n), self.ld)
        return h[: self.n, self.m:]


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def sync():
    torch.cuda.synchronize()
