"""Gated Equihash engine (include/nx_equihash.h).  The reference node contains no Equihash, so there is no
interface to mirror and nothing to be bit-exact against: (n, k) and the 8-byte personalization tag are inputs."""
from __future__ import annotations

import ctypes as C

from ._lib import check, inb, lib, outb


def solution_bytes(n: int, k: int) -> int:
    return lib.nx_equihash_solution_bytes(n, k)


class Solver:
    def __init__(self, n: int, k: int, tag: bytes, device: int = 0):
        assert len(tag) == 8
        h = C.c_void_p()
        check(lib.nx_equihash_create(device, n, k, inb(tag), C.byref(h)))
        self._h, self.n, self.k, self.tag = h, n, k, tag
        self.solution_bytes = solution_bytes(n, k)

    def solve(self, header: bytes, max_solutions: int = 64):
        out = outb(self.solution_bytes * max_solutions)
        cnt = C.c_int()
        check(lib.nx_equihash_solve(self._h, inb(header), len(header), out, max_solutions, C.byref(cnt)))
        raw = bytes(out)
        return [raw[i * self.solution_bytes:(i + 1) * self.solution_bytes] for i in range(min(cnt.value, max_solutions))]

    def validate(self, header: bytes, solutions) -> list:
        solutions = list(solutions)
        if not solutions:
            return []
        v = outb(len(solutions))
        check(lib.nx_equihash_validate(self._h, inb(header), len(header), inb(b"".join(solutions)), len(solutions), v))
        return [bool(v[i]) for i in range(len(solutions))]

    @property
    def last_dropped(self) -> int:
        return lib.nx_equihash_last_dropped(self._h)

    @property
    def last_candidates(self) -> int:
        return lib.nx_equihash_last_candidates(self._h)

    def last_timing(self):
        ms, n = C.c_float(), C.c_int()
        lib.nx_equihash_last_timing(self._h, C.byref(ms), C.byref(n))
        return ms.value, n.value

    def close(self):
        if getattr(self, "_h", None):
            lib.nx_equihash_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
