"""Gated Equihash engine.  NOT a parity test against the reference (it has no Equihash: SURVEY.md §0 Fact 1) -
a self-consistency test: the GPU solver's solution set equals that of an independent, exhaustive CPU
implementation of the published algorithm (oracle/nx_equihash_oracle.c) on fixed headers, every GPU solution
passes the CPU validator and the GPU validator, and tampered solutions are rejected by both."""
import hashlib

import pytest

TAG = b"NodexaPW"  # an arbitrary 8-byte tag: the chain defines none


def _header(nonce: int, length: int = 140) -> bytes:
    base = hashlib.sha512(b"nodexa-equihash-header").digest() * 3
    return base[:length - 32] + nonce.to_bytes(32, "little")


def test_oracle_blake2b_matches_hashlib(oracle):
    d = bytes(range(256)) * 2
    for n, k in ((200, 9), (144, 5), (96, 5)):
        pers = TAG + n.to_bytes(4, "little") + k.to_bytes(4, "little")
        for ln in (0, 1, 127, 128, 129, 140, 144, 300):
            for outlen in (32, (512 // n) * n // 8, 64):
                assert oracle.blake2b(d[:ln], pers, outlen) == hashlib.blake2b(d[:ln], digest_size=outlen, person=pers).digest()


def test_oracle_solver_and_validator_agree(oracle):
    total = 0
    for n, k in ((48, 5), (96, 5)):
        for nonce in range(4):
            hdr = _header(nonce)
            sols = oracle.equihash_solve(n, k, TAG, hdr)
            assert sols == sorted(set(sols))
            for s in sols:
                assert len(s) == oracle.equihash_solution_bytes(n, k)
                assert oracle.equihash_validate(n, k, TAG, hdr, s)
                idx = oracle.equihash_decode(n, k, s)
                assert len(set(idx)) == 1 << k and idx[0] < idx[1 << (k - 1)]
                assert not oracle.equihash_validate(n, k, TAG, _header(nonce + 100), s)  # bound to its header
                bad = bytearray(s)
                bad[len(bad) // 2] ^= 0x04
                assert not oracle.equihash_validate(n, k, TAG, hdr, bytes(bad))
                swapped = oracle.equihash_decode(n, k, s)
                swapped[0], swapped[1] = swapped[1], swapped[0]  # violates the ordering rule
                import ctypes as C
                enc = (C.c_uint8 * len(s))()
                oracle.L.nxeq_encode(n, k, (C.c_uint32 * len(swapped))(*swapped), enc)
                assert not oracle.equihash_validate(n, k, TAG, hdr, bytes(enc))
            total += len(sols)
    assert total > 0


@pytest.mark.gpu
@pytest.mark.parametrize("n,k,nonces", [(48, 5, 12), (96, 5, 6), (80, 4, 6), (200, 9, 1)])
def test_gpu_solution_set_equals_cpu_oracle(nx, gpu, oracle, n, k, nonces):
    found = 0
    with nx.equihash.Solver(n, k, TAG) as s:
        assert s.solution_bytes == oracle.equihash_solution_bytes(n, k)
        for nonce in range(nonces):
            hdr = _header(nonce)
            got = s.solve(hdr)
            want = oracle.equihash_solve(n, k, TAG, hdr)
            assert got == want, (n, k, nonce, len(got), len(want))
            assert s.last_dropped == 0
            assert s.validate(hdr, got) == [True] * len(got)
            found += len(got)
    assert found > 0 or n == 48


@pytest.mark.gpu
@pytest.mark.parametrize("n,k", [(200, 9), (144, 5)])
def test_gpu_full_sized_parameters_validate(nx, gpu, oracle, n, k):
    """(200,9): 2^21 strings, 9 rounds; (144,5): 2^25 strings, 5 rounds.  Every GPU solution must pass the independent
    CPU validator and the GPU validator; tampered, re-ordered and foreign-header solutions must be rejected by both."""
    total = 0
    leaves = 1 << k
    with nx.equihash.Solver(n, k, TAG) as s:
        for nonce in range(6):
            hdr = _header(nonce)
            sols = s.solve(hdr)
            total += len(sols)
            for sol in sols:
                assert oracle.equihash_validate(n, k, TAG, hdr, sol)
            assert s.validate(hdr, sols) == [True] * len(sols)
            if sols:
                bad = bytearray(sols[0]); bad[50] ^= 0x20
                idx = oracle.equihash_decode(n, k, sols[0])
                idx[0], idx[leaves // 2] = idx[leaves // 2], idx[0]
                import ctypes as C
                enc = (C.c_uint8 * len(sols[0]))()
                oracle.L.nxeq_encode(n, k, (C.c_uint32 * leaves)(*idx), enc)
                assert s.validate(hdr, [bytes(bad), bytes(enc), sols[0]]) == [False, False, True]
                assert s.validate(_header(nonce + 50), [sols[0]]) == [False]
                assert not oracle.equihash_validate(n, k, TAG, hdr, bytes(enc))
        ms, launches = s.last_timing()
        assert ms > 0 and launches in (1 + k, 2 + k)
    assert total >= 4  # ~1.9 solutions per nonce are expected


def test_equihash_rejects_unsupported_parameters(nx):
    import ctypes as C
    h = C.c_void_p()
    for n, k in ((200, 8), (100, 4), (64, 1), (512, 15), (0, 0)):
        assert nx.lib.nx_equihash_create(0, n, k, (C.c_uint8 * 8)(*TAG), C.byref(h)) == -1 and not h.value
        assert nx.lib.nx_equihash_solution_bytes(n, k) == 0
    assert nx.lib.nx_equihash_solution_bytes(200, 9) == 1344 and nx.lib.nx_equihash_solution_bytes(144, 5) == 100


@pytest.mark.gpu
@pytest.mark.parametrize("n,k,count", [(96, 5, 9), (200, 9, 12)])
def test_batch_of_headers_equals_one_at_a_time(nx, gpu, n, k, count):
    """nx_equihash_solve_batch keeps two solves in flight on two sets of device buffers: per header it must return exactly
    what nx_equihash_solve returns (same solutions, same order), for odd and even batch lengths, a batch of one, an empty
    batch, and when the output array is too small for everything."""
    hdrs = [_header(1000 + i) for i in range(count)]
    with nx.equihash.Solver(n, k, TAG) as s:
        single = [s.solve(h) for h in hdrs]
        assert any(single), "no header of the batch has a solution: pick other nonces"
        assert s.solve_batch(hdrs) == single
        assert s.last_dropped == 0 and s.last_wall_ms > 0
        assert s.solve_batch(hdrs[:count - 1]) == single[:count - 1]
        assert s.solve_batch(hdrs[:1]) == single[:1]
        assert s.solve_batch([]) == []
        assert [s.solve(h) for h in hdrs[:3]] == single[:3]          # the single-header path still works afterwards
        total = sum(len(x) for x in single)
        if total >= 2:
            got = s.solve_batch(hdrs, max_solutions_total=total - 1)  # one slot short: the last solution is cut, counts are not
            assert sum(len(x) for x in got) == total - 1
            flat = [y for x in single for y in x]
            assert [y for x in got for y in x] == flat[:total - 1]
        with pytest.raises(ValueError):
            s.solve_batch([hdrs[0], hdrs[1] + b"\0\0\0\0"])
