"""Minimal device-memory helper for the GPU tests of the device-resident entry points (cudart through ctypes)."""
import ctypes as C

import numpy as np

_rt = None


def rt():
    global _rt
    if _rt is None:
        _rt = C.CDLL("libcudart.so")
    return _rt


GUARD = 256  # canary bytes in front of and behind every buffer (compute-sanitizer is not available on the GPU pool)


class DevBuf:
    """Device buffer with a poisoned body (an output the kernel did not write shows as 0xCD) and canary zones on both
    sides that every get() checks: a kernel writing outside its output array fails the test that used the buffer."""

    def __init__(self, nbytes):
        self.base = C.c_void_p()
        self.nbytes = int(nbytes)
        self.body = (max(self.nbytes, 4) + 255) // 256 * 256
        assert rt().cudaMalloc(C.byref(self.base), C.c_size_t(self.body + 2 * GUARD)) == 0
        assert rt().cudaMemset(self.base, 0xA5, C.c_size_t(self.body + 2 * GUARD)) == 0
        self.p = C.c_void_p(self.base.value + GUARD)
        assert rt().cudaMemset(self.p, 0xCD, C.c_size_t(max(self.nbytes, 4))) == 0  # poison: unwritten output shows

    @property
    def ptr(self):
        return self.p.value

    def check_guards(self):
        g = np.zeros(GUARD, np.uint8)
        assert rt().cudaDeviceSynchronize() == 0
        for off, what in ((0, "in front of"), (GUARD + max(self.nbytes, 4), "behind")):
            n = GUARD if off == 0 else self.body + GUARD - max(self.nbytes, 4)
            g = np.zeros(n, np.uint8)
            assert rt().cudaMemcpy(C.c_void_p(g.ctypes.data), C.c_void_p(self.base.value + off), C.c_size_t(n), 2) == 0
            assert np.all(g == 0xA5), f"a kernel wrote {what} a {self.nbytes}-byte device buffer"

    @classmethod
    def of(cls, arr):
        arr = np.ascontiguousarray(arr)
        b = cls(arr.nbytes)
        assert rt().cudaMemcpy(b.p, C.c_void_p(arr.ctypes.data), C.c_size_t(arr.nbytes), 1) == 0
        return b

    def get(self, dtype, shape):
        out = np.zeros(shape, dtype)
        assert out.nbytes <= max(self.nbytes, 4)
        self.check_guards()
        assert rt().cudaMemcpy(C.c_void_p(out.ctypes.data), self.p, C.c_size_t(out.nbytes), 2) == 0
        return out

    def free(self):
        if getattr(self, "base", None):
            rt().cudaFree(self.base)
            self.base = self.p = C.c_void_p()

    __del__ = free
