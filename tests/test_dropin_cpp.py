"""The C++ drop-in surface (include/nx/ethash_compat.hpp behind the reference's include paths):
on the CPU it must compile as C++11 (and the C header as C99) and link against libnxkawpow.so;
on the GPU the test program — the reference's kawpow_tests.cpp cases, without Boost — must pass."""
import os
import subprocess

import pytest
from conftest import ROOT, golden


def _gen_header(path):
    k = golden("kawpow_kats.json")
    m = golden("misc.json")
    c = m["ethash_hash"][1]  # epoch 0, nonce 0x123456789abcdef0
    assert c["epoch"] == 0
    lines = ["struct kat { int block_number; const char* header; const char* nonce; const char* mix; const char* final; };",
             "static const kat kats[] = {"]
    for t in k["hash_cases"]:
        lines.append(f'  {{{t["height"]}, "{t["header"]}", "{t["nonce"]:016x}", "{t["mix"]}", "{t["final"]}"}},')
    lines.append("};")
    t = k["hash_30000"]
    lines.append(f'static const kat kat30000 = {{{t["height"]}, "{t["header"]}", "{t["nonce"]:016x}", "{t["mix"]}", "{t["final"]}"}};')
    lines.append(f'static const kat classic_case = {{0, "{c["header"]}", "{c["nonce"]:016x}", "{c["mix"]}", "{c["final"]}"}};')
    lines.append("static const unsigned l1_expected[] = {" + ", ".join(f"{x}u" for x in k["l1_epoch0_first20"]) + "};")
    open(path, "w").write("\n".join(lines) + "\n")


def _build(tmp_path):
    _gen_header(tmp_path / "kats_generated.hpp")
    exe = tmp_path / "dropin_test"
    libdir = os.path.join(ROOT, "nodexa_chain_core_b200")
    subprocess.check_call(["g++", "-std=c++11", "-O1", "-Wall", "-Wextra", "-pthread", "-I", os.path.join(ROOT, "include"), "-I", str(tmp_path),
                           os.path.join(ROOT, "tests", "cpp", "kawpow_dropin_test.cpp"), "-o", str(exe),
                           "-L", libdir, "-lnxkawpow", f"-Wl,-rpath,{libdir}"])
    return exe


def test_headers_compile_and_link(tmp_path):
    _build(tmp_path)
    csrc = tmp_path / "c_view.c"
    csrc.write_text('#include <crypto/ethash/include/ethash/ethash.h>\n#include <crypto/ethash/include/ethash/keccak.h>\n'
                    '#include <nx_kawpow.h>\n'
                    'int main(void) { union ethash_hash256 h = ethash_calculate_epoch_seed(1); '
                    'return (h.bytes[0] == 0x29 && ethash_calculate_light_cache_num_items(0) == 262139 && '
                    'sizeof(struct ethash_result) == 64 && sizeof(nx_search_result) == 80) ? 0 : 1; }\n')
    libdir = os.path.join(ROOT, "nodexa_chain_core_b200")
    exe = tmp_path / "c_view"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-I", os.path.join(ROOT, "include"), str(csrc), "-o", str(exe),
                           "-L", libdir, "-lnxkawpow", f"-Wl,-rpath,{libdir}"])
    assert subprocess.run([str(exe)]).returncode == 0  # host-only calls work without a GPU


@pytest.mark.gpu
def test_reference_kawpow_test_cases_through_cpp_dropin(tmp_path, gpu):
    exe = _build(tmp_path)
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "ALL OK" in out.stdout, out.stdout[-3000:] + out.stderr[-2000:]


def _build_c_example(tmp_path):
    libdir = os.path.join(ROOT, "nodexa_chain_core_b200")
    exe = tmp_path / "mine_and_verify"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "mine_and_verify.c"), "-o", str(exe), "-L", libdir, "-lnxkawpow",
                           f"-Wl,-rpath,{libdir}"])
    return exe


def test_c_example_compiles(tmp_path):
    _build_c_example(tmp_path)


@pytest.mark.gpu
def test_c_example_mines_and_verifies(tmp_path, gpu):
    """examples/mine_and_verify.c: the node-side flow in plain C99 over the C ABI."""
    out = subprocess.run([str(_build_c_example(tmp_path))], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and out.stdout.strip().endswith("OK"), out.stdout[-1000:] + out.stderr[-1000:]


REF_TESTS = os.path.join(ROOT, "oracle", "_ref", "kawpow_tests_reference")


def test_reference_test_file_builds_against_dropin_headers():
    """The reference's own src/test/kawpow_tests.cpp, unmodified, compiles against include/crypto/ethash/... and links with
    libnxkawpow.so (oracle/Makefile `ref_tests`).  Only where /root/reference exists; the binary then travels to the GPU box."""
    if not os.path.isdir("/root/reference/src/test"):
        pytest.skip("reference tree not present here")
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "ref_tests"])
    assert os.path.exists(REF_TESTS)


@pytest.mark.gpu
def test_reference_test_file_passes_on_the_gpu(gpu):
    """kawpow_l1_cache, kawpow_hash_empty (1000 heights), kawpow_hash_30000, kawpow_hash_and_verify (13 vectors + negatives),
    kawpow_search - the reference's test cases, the reference's source file, this repository's implementation."""
    if not os.path.exists(REF_TESTS):
        pytest.skip("oracle/_ref/kawpow_tests_reference was not built (needs /root/reference at build time)")
    out = subprocess.run([REF_TESTS], capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "*** 5 test cases" in out.stdout and " 0 failures" in out.stdout, out.stdout[-500:]
