"""nx_dag_checksum_noalloc: the DAG checksum for callers that must not allocate.

Some users of this library cannot call cudaMalloc: a watchdog thread that checks the DAG while a search
is running (an allocation would stall every stream of that GPU), or a plug-in inside an application that
owns all of the GPU's memory.  For them the library carries a fixed 16.25 KiB scratch array; this test
drives that entry point through the C ABI and compares it, bit for bit, with nx_dag_checksum (which
uses a buffer allocated by the epoch context).  Calls are made back to back on purpose: the array
keeps the previous call's contents, and a smaller input after a larger one must not see them.

Collected last (file name) so that a failure here cannot hide any other test under `pytest -x`.
"""
import ctypes as C

import numpy as np
import pytest
from conftest import golden

pytestmark = pytest.mark.gpu

NX_NOALLOC_MAX_BYTES = 8 << 30


def _noalloc(nx, ptr, nbytes):
    out = C.c_uint64(0xDEADBEEF)
    rc = nx.lib.nx_dag_checksum_noalloc(C.c_void_p(ptr), nbytes, None, C.byref(out))
    return rc, out.value


def _weighted_sum(buf: bytes) -> int:
    w = np.frombuffer(buf, dtype="<u8")
    i = np.arange(w.size, dtype=np.uint64)
    with np.errstate(over="ignore"):
        return int((w * (i * np.uint64(2) + np.uint64(1))).sum(dtype=np.uint64))


def test_noalloc_checksum_matches_allocating_entry_point_twice_in_a_row(nx, ctx_of):
    ctx = ctx_of(0)
    want = ctx.checksum()  # existing entry point (context-owned, cudaMalloc'ed partials)
    assert want == golden("dag_digest.json")["0"]["weighted_sum"]

    rc, got1 = _noalloc(nx, ctx.dag_ptr, ctx.dag_bytes)  # first call on this device: does the one-time set-up
    assert rc == 0, nx.lib.nx_last_error()
    assert got1 == want

    # a smaller input right after the 1 GiB one: stale partials from the first call must not leak in
    small = 4096 * 256
    rc, got_small = _noalloc(nx, ctx.dag_ptr, small)
    assert rc == 0, nx.lib.nx_last_error()
    assert got_small == golden("dag_items.json")["0"]["first4096_weighted_sum"]
    assert got_small == _weighted_sum(ctx.read_items(0, 4096))

    rc, got2 = _noalloc(nx, ctx.dag_ptr, ctx.dag_bytes)  # second full call, array still holds the last one
    assert rc == 0, nx.lib.nx_last_error()
    assert got2 == want
    assert ctx.checksum() == want  # and the allocating path is undisturbed


def test_noalloc_checksum_ragged_sizes(nx, ctx_of):
    """Sizes that are not a multiple of the 4 MiB chunk, a single word, and the empty input."""
    ctx = ctx_of(0)
    for nbytes in (8, 264, (4 << 20) - 8, (4 << 20) + 8, 3 * (4 << 20) + 1024):
        rc, got = _noalloc(nx, ctx.dag_ptr, nbytes)
        assert rc == 0, nx.lib.nx_last_error()
        items = (nbytes + 255) // 256
        assert got == _weighted_sum(ctx.read_items(0, items)[:nbytes]), nbytes
    rc, got = _noalloc(nx, ctx.dag_ptr, 0)
    assert (rc, got) == (0, 0)


def test_noalloc_checksum_rejects_what_it_cannot_serve(nx, ctx_of):
    ctx = ctx_of(0)
    rc, _ = _noalloc(nx, ctx.dag_ptr, NX_NOALLOC_MAX_BYTES + 8)  # beyond the fixed scratch: refused, not truncated
    assert rc == -1
    msg = nx.lib.nx_last_error().decode()
    assert "exceed" in msg and "nx_dag_checksum" in msg
    assert _noalloc(nx, ctx.dag_ptr, 12)[0] == -1       # not whole 64-bit words
    assert _noalloc(nx, ctx.dag_ptr + 4, 8)[0] == -1    # misaligned
    assert _noalloc(nx, 0, 8)[0] == -1                  # null input
    host = (C.c_uint64 * 4)()
    assert _noalloc(nx, C.addressof(host), 32)[0] < 0   # host memory is not accepted
    rc = nx.lib.nx_dag_checksum_noalloc(C.c_void_p(ctx.dag_ptr), 8, None, None)
    assert rc == -1                                      # null output
    # a range that leaves the allocation is refused up front instead of faulting inside the kernel
    assert _noalloc(nx, ctx.dag_ptr, ctx.dag_padded_bytes + (64 << 20))[0] == -1
    assert "past the end" in nx.lib.nx_last_error().decode()
    assert _noalloc(nx, ctx.dag_ptr + (8 << 20), ctx.dag_padded_bytes)[0] == -1
    rc, got = _noalloc(nx, ctx.dag_ptr, ctx.dag_bytes)  # and the device is still healthy
    assert rc == 0 and got == golden("dag_digest.json")["0"]["weighted_sum"]
