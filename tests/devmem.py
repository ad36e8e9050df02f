"""Minimal device-memory helper for the GPU tests of the device-resident entry points (cudart through ctypes)."""
import ctypes as C

import numpy as np

_rt = None


def rt():
    global _rt
    if _rt is None:
        _rt = C.CDLL("libcudart.so")
    return _rt


class DevBuf:
    def __init__(self, nbytes):
        self.p = C.c_void_p()
        self.nbytes = int(nbytes)
        assert rt().cudaMalloc(C.byref(self.p), C.c_size_t(max(self.nbytes, 4))) == 0
        assert rt().cudaMemset(self.p, 0xCD, C.c_size_t(max(self.nbytes, 4))) == 0  # poison: unwritten output shows

    @property
    def ptr(self):
        return self.p.value

    @classmethod
    def of(cls, arr):
        arr = np.ascontiguousarray(arr)
        b = cls(arr.nbytes)
        assert rt().cudaMemcpy(b.p, C.c_void_p(arr.ctypes.data), C.c_size_t(arr.nbytes), 1) == 0
        return b

    def get(self, dtype, shape):
        out = np.zeros(shape, dtype)
        assert out.nbytes <= max(self.nbytes, 4)
        assert rt().cudaDeviceSynchronize() == 0
        assert rt().cudaMemcpy(C.c_void_p(out.ctypes.data), self.p, C.c_size_t(out.nbytes), 2) == 0
        return out

    def free(self):
        if self.p:
            rt().cudaFree(self.p)
            self.p = C.c_void_p()

    __del__ = free
