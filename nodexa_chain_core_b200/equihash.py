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

    def solve_batch(self, headers, max_solutions_total: int = 0) -> list:
        """Solutions of several headers of equal length, two solves in flight on the GPU: [[solution, ...], ...] in
        header order.  last_wall_ms is the host time of the whole call."""
        headers = list(headers)
        if not headers:
            return []
        hl = len(headers[0])
        if any(len(h) != hl for h in headers):
            raise ValueError("headers of one batch must have the same length")
        cap = max_solutions_total or 16 * len(headers)
        out = outb(self.solution_bytes * cap)
        counts = (C.c_int * len(headers))()
        ms = C.c_float()
        check(lib.nx_equihash_solve_batch(self._h, inb(b"".join(headers)), hl, len(headers), out, cap, counts, C.byref(ms)))
        self.last_wall_ms = ms.value
        raw, res, pos = bytes(out), [], 0
        for c in counts:
            take = max(0, min(c, cap - pos))
            res.append([raw[(pos + j) * self.solution_bytes:(pos + j + 1) * self.solution_bytes] for j in range(take)])
            pos += take
        return res

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
