"""External miner over the node's JSON-RPC - the route real KawPow miners take today.

The node hands out work through `getblocktemplate`: next to the usual fields it adds `pprpcheader` (hex of
GetKAWPOWHeaderHash(), i.e. the 32-byte header hash in the byte order progpow::hash takes) and `pprpcepoch`
(reference src/rpc/mining.cpp:722-737; needs `-miningaddress`), and it remembers the template under that header hash.
`target` is the 256-bit boundary as big-endian hex (:697), `height` the block height (:716).  A solution goes back as
`pprpcsb "<header_hash>" "<mix_hash>" "<nonce hex>"` (:842-930), which re-hashes on the node and runs ProcessNewBlock.

This module is that client on top of nx_miner_* (every visible GPU, lowest nonce first): poll, search a window of the
nonce space, submit on a hit, start over when the template changes.  Transport is a callable, so tests can stand in for
the node; `http_rpc()` builds the real one (HTTP POST with basic auth, like clore-cli).
"""
from __future__ import annotations

import base64
import json
import time
import urllib.error
import urllib.request
from dataclasses import dataclass, field

from .miner import Miner


def http_rpc(url: str, user: str = "", password: str = "", timeout: float = 30.0):
    """JSON-RPC 1.0 over HTTP as the node's RPC server speaks it; returns call(method, params) -> result."""
    auth = base64.b64encode(f"{user}:{password}".encode()).decode()
    counter = [0]

    def call(method: str, params: list):
        counter[0] += 1
        body = json.dumps({"jsonrpc": "1.0", "id": counter[0], "method": method, "params": params}).encode()
        req = urllib.request.Request(url, data=body, headers={"Content-Type": "application/json", "Authorization": "Basic " + auth})
        try:
            with urllib.request.urlopen(req, timeout=timeout) as resp:  # noqa: S310 - the URL is the operator's own node
                reply = json.loads(resp.read().decode())
        except urllib.error.HTTPError as e:
            # the node answers RPC-level failures with HTTP 500/404 and the JSON-RPC error in the body (httprpc.cpp)
            try:
                reply = json.loads(e.read().decode())
            except ValueError:
                raise e from None  # 401 and friends carry no JSON: let the transport error through
        if reply.get("error"):
            raise RuntimeError(f"RPC {method} failed: {reply['error']}")
        return reply["result"]

    return call


@dataclass
class Work:
    """The part of a block template an external KawPow miner needs."""
    header_hash: bytes      # pprpcheader
    boundary: bytes         # target, 32 big-endian bytes
    height: int
    epoch: int              # pprpcepoch
    next_nonce: int = 0

    @staticmethod
    def from_template(t: dict) -> "Work":
        if "pprpcheader" not in t:
            raise ValueError("getblocktemplate reply has no pprpcheader: start the node with -miningaddress=<address> "
                             "and make sure the KawPow activation time has passed")
        header = bytes.fromhex(t["pprpcheader"])
        target = bytes.fromhex(t["target"].rjust(64, "0"))
        if len(header) != 32 or len(target) != 32:
            raise ValueError("malformed pprpcheader / target")
        height = int(t["height"])
        epoch = int(t.get("pprpcepoch", height // 7500))
        if epoch != height // 7500:
            raise ValueError(f"pprpcepoch {epoch} does not match height {height}")
        return Work(header, target, height, epoch)


@dataclass
class RpcMiner:
    rpc: object                               # call(method, params) -> result
    miner: Miner = None                       # default: all visible GPUs
    window: int = 1 << 28                     # nonces per search call (1.4 s on one B200)
    start_nonce: int = 0                      # pools give each rig its own range; solo mining starts anywhere
    lookahead_blocks: int = 60                # prebuild the next epoch this many blocks ahead
    template_request: dict = field(default_factory=lambda: {"rules": ["segwit"]})
    work: Work = None
    submitted: list = field(default_factory=list)
    hashes: int = 0
    seconds: float = 0.0

    def __post_init__(self):
        if self.miner is None:
            self.miner = Miner()

    def poll(self) -> bool:
        """Fetch the current template; True when it differs from the one being worked on (new tip, new transactions,
        new time) - the nonce range then starts over, as the header hash commits to all of those."""
        w = Work.from_template(self.rpc("getblocktemplate", [self.template_request]))
        if self.work is not None and w.header_hash == self.work.header_hash:
            return False
        w.next_nonce = self.start_nonce
        self.work = w
        return True

    def step(self):
        """One search window on the current work.  Returns the node's answer to pprpcsb when a solution was found and
        submitted (True / None mean accepted, a string is the node's reason), else None."""
        if self.work is None:
            self.poll()
        w = self.work
        self.miner.prepare_async(w.height, self.lookahead_blocks)   # next period's program, next epoch's DAGs
        count = min(self.window, (1 << 64) - w.next_nonce)
        sr = self.miner.search(w.height, w.header_hash, w.boundary, w.next_nonce, count)
        n, s = self.miner.last_stats()
        self.hashes += n
        self.seconds += s
        if not sr.solution_found:
            w.next_nonce = (w.next_nonce + count) % (1 << 64)
            return None
        # pprpcsb takes the mix as the node prints it (hex of the hash256 bytes) and the nonce as hex (ParseUInt64 base 16)
        answer = self.rpc("pprpcsb", [w.header_hash.hex(), sr.mix_hash.hex(), "%x" % sr.nonce])
        self.submitted.append({"height": w.height, "header_hash": w.header_hash.hex(), "nonce": sr.nonce,
                               "mix_hash": sr.mix_hash.hex(), "final_hash": sr.final_hash.hex(), "answer": answer})
        self.work = None  # solved or stale either way: the next step asks for a fresh template
        return True if answer is None else answer

    def run(self, max_blocks: int | None = None, max_seconds: float | None = None, poll_every: int = 4):
        """Mine until max_blocks solutions were submitted or max_seconds passed; re-polls the template every
        `poll_every` windows so that a new tip is noticed within a few seconds."""
        t0 = time.perf_counter()
        windows = 0
        while (max_blocks is None or len(self.submitted) < max_blocks) and \
                (max_seconds is None or time.perf_counter() - t0 < max_seconds):
            if self.work is None or windows % poll_every == 0:
                self.poll()
            self.step()
            windows += 1
        return self.submitted

    @property
    def hashrate(self) -> float:
        """H/s over the search calls so far (what getmininginfo calls hashespersec)."""
        return self.hashes / self.seconds if self.seconds else 0.0

    def close(self):
        self.miner.close()
