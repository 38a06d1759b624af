"""The external-miner client (nodexa_chain_core_b200/rpc_miner.py) against a mock of the node's mining RPCs served over
HTTP (tests/mock_node.py).  On the CPU the GPU miner is replaced by the oracle's search, which exercises the whole
protocol (template fields, hex byte orders, nonce ranges, template changes, rejected submissions, HTTP errors); the GPU
test then runs the real nx_miner through the same loop."""
import urllib.error

import pytest
from mock_node import MockNode

from nodexa_chain_core_b200 import node
from nodexa_chain_core_b200.progpow import SearchResult
from nodexa_chain_core_b200.rpc_miner import RpcMiner, Work, http_rpc


class OracleMiner:
    """Same surface as nodexa_chain_core_b200.miner.Miner, backed by the CPU oracle (small windows only)."""

    def __init__(self, oracle):
        self.oracle, self.prepared, self.searched = oracle, [], 0

    def prepare_async(self, height, lookahead):
        self.prepared.append((height, lookahead))

    def search(self, height, header_hash, boundary, start, count):
        found, nonce, fin, mix = self.oracle.kawpow_search(height // 7500, height, header_hash, boundary, start, count)
        self.searched = count
        return SearchResult(found, nonce, fin, mix)

    def last_stats(self):
        return self.searched, 0.5

    def close(self):
        pass


def test_work_from_template():
    t = {"pprpcheader": "ab" * 32, "pprpcepoch": 1, "target": "ff" * 31, "height": 7500}
    w = Work.from_template(t)
    assert (w.header_hash, w.boundary, w.height, w.epoch) == (bytes.fromhex("ab" * 32), bytes.fromhex("00" + "ff" * 31), 7500, 1)
    with pytest.raises(ValueError, match="miningaddress"):
        Work.from_template({"target": "ff", "height": 1})
    with pytest.raises(ValueError, match="pprpcepoch"):
        Work.from_template({**t, "pprpcepoch": 0})
    with pytest.raises(ValueError):
        Work.from_template({**t, "pprpcheader": "abcd"})


def test_protocol_round_trip_over_http_with_the_oracle_as_miner(oracle):
    mock = MockNode(oracle, first_height=100)
    url = mock.serve()
    try:
        with pytest.raises(urllib.error.HTTPError):                      # wrong password: 401, like the node
            http_rpc(url, "rpcuser", "nope")("getmininginfo", [])
        rpc = http_rpc(url, "rpcuser", "rpcpass")
        assert rpc("getmininginfo", []) == {"blocks": 99, "chain": "mock"}
        with pytest.raises(RuntimeError, match="not found"):
            rpc("pprpcsb", ["00" * 32, "00" * 32, "0"])
        m = RpcMiner(rpc, miner=OracleMiner(oracle), window=64, start_nonce=1 << 40)
        assert m.poll() and not m.poll()                                 # same template twice: keep the nonce range
        first = m.work.header_hash
        mock.bump += 1                                                   # "new transactions": the header hash changes
        assert m.poll() and m.work.header_hash != first and m.work.next_nonce == 1 << 40
        got = m.run(max_blocks=2, max_seconds=120)
        assert len(got) == 2 and [g["answer"] for g in got] == [None, None] and len(mock.accepted) == 2
        assert [b.height for b in mock.accepted] == [100, 101] and mock.accepted[1].prev_block == mock.accepted[0].block_hash()
        for g, b in zip(got, mock.accepted):
            assert g["nonce"] == b.nonce64 >= 1 << 40 and bytes.fromhex(g["mix_hash"]) == b.mix_hash[::-1]
            assert bytes.fromhex(g["final_hash"]) == b.block_hash()[::-1]            # the block id is the PoW hash
            ok, target = node.compact_to_boundary(b.bits)
            # lowest nonce of the scanned range: nothing below it qualified
            assert not oracle.kawpow_search(0, b.height, b.kawpow_header_hash(), target, 1 << 40, g["nonce"] - (1 << 40))[0]
        assert m.miner.prepared and m.miner.prepared[0] == (100, 60) and m.hashrate > 0
        # a wrong mix is turned away by the node and reported, not swallowed
        m.poll()
        w = m.work
        found, nonce, fin, mix = oracle.kawpow_search(0, w.height, w.header_hash, w.boundary, 0, 4096)
        assert found and rpc("pprpcsb", [w.header_hash.hex(), "ab" * 32, "%x" % nonce]) == "invalid-mix-hash"
        with pytest.raises(RuntimeError, match="does not solve"):
            miss = next(n for n in range(64) if oracle.kawpow_hash(0, w.height, w.header_hash, n)[0] > w.boundary)
            rpc("pprpcsb", [w.header_hash.hex(), mix.hex(), "%x" % miss])
        assert "getblocktemplate" in mock.calls and mock.calls.count("pprpcsb") >= 4
    finally:
        mock.stop()


@pytest.mark.gpu
def test_gpu_miner_mines_a_chain_through_the_rpc_protocol(nx, oracle, gpu):
    """The real thing: nx_miner on the GPU, templates and submissions over HTTP, the mock node judging every block with
    the CPU oracle - five linked blocks across a period switch, each the lowest nonce of its range."""
    mock = MockNode(oracle, first_height=7495, bits=0x1f00ffff)     # 1 nonce in 65536; crosses into epoch 1
    url = mock.serve()
    try:
        with nx.Miner([gpu]) as gm:
            gm.set_batch(1 << 16)
            m = RpcMiner(http_rpc(url, "rpcuser", "rpcpass"), miner=gm, window=1 << 22, lookahead_blocks=6)
            got = m.run(max_blocks=7, max_seconds=300)
        assert [g["answer"] for g in got] == [None] * 7 and not mock.rejected
        assert [b.height for b in mock.accepted] == list(range(7495, 7502))
        for a, b in zip(mock.accepted, mock.accepted[1:]):
            assert b.prev_block == a.block_hash()
        assert m.hashrate > 1e6
    finally:
        mock.stop()
