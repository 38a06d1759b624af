"""The product's host code (csrc/nx_host.cpp, nx_codegen.cpp, nx_lightcache_avx512.cpp ...) rebuilt with
AddressSanitizer + UndefinedBehaviorSanitizer and driven by tests/cpp/host_sanitize_test.cpp: no report, and every
value equal to what the regular library and the oracle give ("race detection / sanitizers";
compute-sanitizer for the device side is closed on this pool).  No GPU needed."""
import ctypes as C
import os
import shutil
import struct
import subprocess

import pytest

from nodexa_chain_core_b200 import ethash
from nodexa_chain_core_b200._lib import inb, lib, outb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "nodexa_chain_core_b200", "csrc")
FLAGS = ["-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-fno-omit-frame-pointer",
         "-I", os.path.join(ROOT, "include"), "-I", SRC, "-I", "/usr/local/cuda/include"]


@pytest.fixture(scope="module")
def san(tmp_path_factory):
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    d = tmp_path_factory.mktemp("san")
    plain = [os.path.join(ROOT, "tests", "cpp", "host_sanitize_test.cpp")] + [
        os.path.join(SRC, f) for f in ("nx_host.cpp", "nx_codegen.cpp", "nx_kernel_src.cpp", "nx_cpu_dispatch.cpp")]
    subprocess.check_call(["g++", *FLAGS, "-c", *plain], cwd=d)
    subprocess.check_call(["g++", *FLAGS, "-mavx512f", "-c", os.path.join(SRC, "nx_lightcache_avx512.cpp")], cwd=d)
    objs = [f for f in os.listdir(d) if f.endswith(".o")]
    subprocess.check_call(["g++", "-fsanitize=address,undefined", *objs, "-o", "host_san", "-lpthread"], cwd=d)
    cache = d / "cache"
    cache.mkdir()
    env = dict(os.environ, NX_CACHE_DIR=str(cache), ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1")
    run = subprocess.run([str(d / "host_san")], capture_output=True, text=True, env=env, timeout=600)
    return run


def _kv(run):
    return dict(line.split("=", 1) for line in run.stdout.splitlines() if "=" in line)


def test_no_sanitizer_report(san):
    assert san.returncode == 0, san.stderr[-3000:]
    assert "runtime error" not in san.stderr and "AddressSanitizer" not in san.stderr and "LeakSanitizer" not in san.stderr, san.stderr[-3000:]
    assert _kv(san).get("done") == "1"


def test_values_match_library_and_oracle(san, oracle):
    kv = _kv(san)
    msg = bytes((i * 7 + 1) & 0xff for i in range(300))
    for ln in (0, 1, 31, 32, 64, 71, 72, 73, 135, 136, 137, 144, 200, 272, 300):
        assert kv[f"keccak256_{ln}"] == oracle.keccak256(msg[:ln]).hex() == ethash.keccak256(msg[:ln]).hex()
        assert kv[f"keccak512_{ln}"] == oracle.keccak512(msg[:ln]).hex() == ethash.keccak512(msg[:ln]).hex()
    for e in (0, 1, 2, 22, 448):
        assert kv[f"seed_{e}"] == oracle.epoch_seed(e).hex()
        assert int(kv[f"light_items_{e}"]) == ethash.calculate_light_cache_num_items(e)
        assert int(kv[f"dataset_items_{e}"]) == ethash.calculate_full_dataset_num_items(e)
    assert [int(kv[f"find_epoch_{e}"]) for e in (0, 1, 2, 22)] == [0, 1, 2, 22]
    for ub in (-5, 0, 1, 2, 3, 4, 100, 262144):
        assert int(kv[f"prime_{ub}"]) == oracle.L.nxo_find_largest_prime(ub)
    assert kv.get("light_avx512_1021", kv["light_scalar_1021"]) == kv["light_scalar_1021"]
    assert (kv["cache_load"], kv["cache_same"], kv["cache_wrong_size"], kv["cache_corrupt"]) == ("1", "1", "0", "0")
    for seed in (0, 1, 2499, 1120000, (1 << 40) + 3):
        arr = ((C.c_uint32 * 6) * 40)()  # the regular library's op list, itself pinned to the oracle in test_codegen.py
        n_ops = lib.nx_kawpow_program_ops(seed, arr, 40)
        ops = [tuple(arr[i]) for i in range(n_ops)]
        flat = b"".join(struct.pack("<6I", *op) for op in ops)
        o = outb(32)
        lib.nx_sha256d(inb(flat), len(flat), o)
        assert kv[f"program_ops_{seed}"] == f"{len(ops)}:{bytes(o).hex()}"
    for bits in (0, 0x00123456, 0x01003456, 0x01123456, 0x02008000, 0x03123456, 0x04123456, 0x04923456, 0x1b0404cb, 0x1f00ffff,
                 0x2000ffff, 0x207fffff, 0x20123456, 0x21123456, 0x22123456, 0x23000001, 0xff123456):
        ok, target = oracle.compact_to_boundary(bits)
        got_ok, got_t = kv[f"compact_{bits:08x}"].split(":")
        assert int(got_ok) == int(ok)
        if ok:
            assert got_t == target.hex()
    n, rec_hex = kv["disk_record"].split(":")
    rec, used = __import__("nodexa_chain_core_b200.node", fromlist=["x"]).DiskBlockIndex.parse(bytes.fromhex(rec_hex))
    assert used == int(n) and (rec.height, rec.file, rec.data_pos, rec.nonce64) == (3360000, 12, 0xfedcba98, 0x0123456789abcdef)
    assert int(kv["disk_fuzz_accepted"]) > 0 and kv["disk_check"] in ("0", "1")
    hh = oracle.kawpow_header_hash(0x30000000, bytes(range(32)), bytes(255 - i for i in range(32)), 1651444217, 0x1b0404cb, 3360000)
    assert kv["header_hash"] == hh.hex() and kv["roundtrip"] == "1"
    mix = bytes((i * 3) & 0xff for i in range(32))
    assert kv["no_verify"] == oracle.kawpow_hash_no_verify(3360000, hh, mix, 0x123456789abcdef0).hex()
