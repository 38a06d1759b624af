"""Product code generator (csrc/nx_codegen.cpp) and kernel text (csrc/kawpow_kernel.cuh), checked
on the CPU: op list vs the oracle's, the generated program executed with emulated lanes vs the
oracle's hashes, and an NVRTC compile for sm_100a (no device needed) with its resource usage."""
import ctypes as C
import os
import re
import shutil
import subprocess

import pytest
from conftest import ROOT


def _program_ops(nx, seed):
    ops = ((C.c_uint32 * 6) * 40)()
    n = nx.lib.nx_kawpow_program_ops(seed, ops, 40)
    return [tuple(ops[i]) for i in range(n)]


def _source(nx, seed, items, skeleton=1):
    n = nx.lib.nx_kawpow_program_source(seed, items, skeleton, None, 0)
    buf = C.create_string_buffer(n + 1)
    nx.lib.nx_kawpow_program_source(seed, items, skeleton, buf, n + 1)
    return buf.value.decode()


@pytest.mark.parametrize("seed", [0, 1, 2, 16, 2499, 10000, 1120000, 2**32 + 5])
def test_program_ops_match_oracle(nx, oracle, seed):
    """Same draws, same order as mix_rng_state + round (progpow.cpp:46-88,191-233)."""
    mine = _program_ops(nx, seed)
    theirs = oracle.kawpow_program(seed)
    assert len(mine) == len(theirs) == 33
    for a, b in zip(mine, theirs):
        kind = a[0]
        assert kind == b[0]
        if kind == 0:    # cache: oracle {0, src, dst, sel}
            assert (a[1], a[3], a[5]) == (b[1], b[2], b[3])
        elif kind == 1:  # math: oracle {1, src1, src2, sel1, dst, sel2}
            assert (a[1], a[2], a[4], a[3], a[5]) == (b[1], b[2], b[3], b[4], b[5])
        else:            # dag: oracle {2, dst, sel}
            assert (a[3], a[5]) == (b[1], b[2])


def test_program_matches_survey_appendix(nx):
    """SURVEY.md appendix A, prog_seed 0: cache[0] src 29 -> dst 18 merge 3 x 6; math[0] r21,r14 op 3 -> dst 31 merge 0."""
    ops = _program_ops(nx, 0)
    k, src, _, dst, _, sel = ops[0]
    assert (k, src, dst, sel % 4, ((sel >> 16) % 31) + 1) == (0, 29, 18, 3, 6)
    k, s1, s2, dst, selm, sel = ops[1]
    assert (k, s1, s2, selm % 11, dst, sel % 4) == (1, 21, 14, 3, 31, 0)
    txt = _source(nx, 0, 4194296, 0)
    assert "m[18] = nx_rotr_c(m[18], 6u, 0u) ^ t0" in txt and "nx_min(m[21], m[14])" in txt


@pytest.mark.parametrize("height", [0, 3, 49 * 3, 7499])
def test_generated_program_on_cpu_vs_oracle(nx, oracle, tmp_path, height):
    """Compile the generated CUDA text as plain C++ and run whole hashes through it."""
    seed = height // 3
    items = oracle.dataset_items(0) // 2
    gen = tmp_path / "gen.h"
    gen.write_text(_source(nx, seed, items, 1))
    so = tmp_path / "emul.so"
    oracle_so = os.path.join(ROOT, "oracle", "libnxoracle.so")
    subprocess.check_call(["g++", "-O1", "-shared", "-fPIC", "-w", "-include", str(gen),
                           os.path.join(ROOT, "tests", "cpu_emul", "kawpow_emul.cpp"), "-o", str(so),
                           oracle_so, f"-Wl,-rpath,{os.path.dirname(oracle_so)}"])
    em = C.CDLL(str(so))
    u8p = C.POINTER(C.c_uint8)
    em.emul_kawpow_hash.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), u8p, C.c_uint64, u8p, u8p]
    ctx = oracle.ctx(0)
    l1 = (C.c_uint32 * 4096)(*oracle.l1(0))
    import random
    rnd = random.Random(height)
    for _ in range(3):
        hd = rnd.randbytes(32)
        nonce = rnd.getrandbits(64)
        f, m = (C.c_uint8 * 32)(), (C.c_uint8 * 32)()
        em.emul_kawpow_hash(C.cast(ctx, C.c_void_p), l1, (C.c_uint8 * 32).from_buffer_copy(hd), nonce, f, m)
        assert (bytes(f), bytes(m)) == oracle.kawpow_hash(0, height, hd, nonce)


def test_nvrtc_compiles_for_sm100a_without_spills(nx, tmp_path):
    """NVRTC needs no device: the period program must build for sm_100a, stay within 64 registers
    (1024 threads per SM) and not spill."""
    buf = C.create_string_buffer(8 << 20)
    n = nx.lib.nx_kawpow_compile_cubin(1120000, 18874358, buf, len(buf))
    assert n > 0, nx.lib.nx_last_error()
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    p = tmp_path / "k.cubin"
    p.write_bytes(buf.raw[:n])
    elf = subprocess.check_output([cuobjdump, "-elf", str(p)], text=True)
    assert re.search(r"arch = sm_100a|EF_CUDA_SM100|sm_100", elf), elf[:400]
    out = subprocess.check_output([cuobjdump, "-res-usage", str(p)], text=True)
    usage = re.findall(r"Function (\w+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", out)
    assert {u[0] for u in usage} == {"nx_kawpow_kernel"}
    for name, reg, stack, shared, local in usage:
        assert int(reg) <= 64, (name, reg)
        assert int(stack) == 0 and int(local) == 0, (name, stack, local)
        assert int(shared) <= 48 * 1024  # static part only; the L1 replicas are dynamic shared memory
