"""Classic Ethash (kernel K7) through the C ABI: bit-exact against the reference-made fixture and the
oracle.  The reference exports this API but nothing in the node calls it."""
import random

import pytest
from conftest import golden

from nodexa_chain_core_b200 import ethash

pytestmark = pytest.mark.gpu


def test_ethash_hash_vs_reference_fixture(ctx_of):
    for c in golden("misc.json")["ethash_hash"]:
        ctx = ctx_of(c["epoch"])
        f, m = ethash.hash(ctx, bytes.fromhex(c["header"]), c["nonce"])
        assert (f.hex(), m.hex()) == (c["final"], c["mix"])
        assert ethash.verify(ctx, bytes.fromhex(c["header"]), m, c["nonce"], f)
        lower = (int.from_bytes(f, "big") - 1).to_bytes(32, "big")
        assert not ethash.verify(ctx, bytes.fromhex(c["header"]), m, c["nonce"], lower)
        bad = bytes([m[0] ^ 1]) + m[1:]
        assert not ethash.verify(ctx, bytes.fromhex(c["header"]), bad, c["nonce"], bytes([0xff]) * 32)


def test_ethash_batch_vs_oracle_live(ctx_of, oracle):
    """Includes the odd last 128-byte DAG item, which KawPow never touches but classic Ethash can."""
    ctx = ctx_of(0)
    rnd = random.Random(9)
    hd = rnd.randbytes(32)
    start = rnd.getrandbits(63)
    n = 300
    got = ethash.hash_batch(ctx, hd, start, n)
    for i in sorted(rnd.sample(range(n), 12)) + [0, n - 1]:
        f, m = oracle.ethash_hash(0, hd, start + i)
        assert got[i * 64:i * 64 + 32] == f and got[i * 64 + 32:i * 64 + 64] == m, i
    assert ethash.hash_batch(ctx, hd, start, 0) == b""
    # the extra half item behind the KawPow range equals the oracle's 128-byte item n-1
    last128 = ctx.read_items(ctx.dag_num_items, 1)[:128]
    n128 = ctx.full_dataset_num_items
    assert last128 == oracle.dataset_item_512(0, 2 * (n128 - 1)) + oracle.dataset_item_512(0, 2 * (n128 - 1) + 1)
