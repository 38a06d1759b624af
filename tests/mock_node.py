"""A stand-in for the node's mining RPCs (reference src/rpc/mining.cpp: getblocktemplate :722-737, pprpcsb :842-930,
getmininginfo) served over real HTTP with basic auth, for tests/test_rpc_miner.py.  Templates are synthetic headers; a
submission is judged by the CPU ORACLE exactly as pprpcsb judges it: the stored template gets the nonce and the mix,
the full hash is recomputed and checked against nBits, and the recomputed mix must equal the submitted one
(CheckBlockHeader, validation.cpp:11652-11662).  An accepted block becomes the tip and the next template builds on it."""
import base64
import hashlib
import json
import threading
from http.server import BaseHTTPRequestHandler, HTTPServer

from nodexa_chain_core_b200 import node


class MockNode:
    def __init__(self, oracle, first_height=100, bits=0x2000ffff, user="rpcuser", password="rpcpass"):
        self.oracle, self.height, self.bits = oracle, first_height, bits
        self.prev = hashlib.sha256(b"mock genesis").digest()
        self.templates = {}          # mapHVNKAWBlockTemplates (mining.cpp:736)
        self.accepted, self.rejected, self.calls = [], [], []
        self.bump = 0                # changes the merkle root, like new mempool transactions would
        self.auth = "Basic " + base64.b64encode(f"{user}:{password}".encode()).decode()
        self.server = None

    # ---- the RPCs ----
    def getblocktemplate(self, _request=None):
        h = node.BlockHeader(0x30000000, self.prev, hashlib.sha256(b"merkle %d %d" % (self.height, self.bump)).digest(),
                             node.MAINNET_KAWPOW_ACTIVATION_TIME + 60 * self.height, self.bits, self.height)
        f = node.pprpc_fields(h)
        self.templates[f["pprpcheader"]] = h
        ok, target = node.compact_to_boundary(self.bits)
        assert ok
        return {"version": h.version, "previousblockhash": self.prev[::-1].hex(), "target": target.hex(), "bits": "%08x" % self.bits,
                "height": self.height, "curtime": h.time, "noncerange": "00000000ffffffff", **f}

    def pprpcsb(self, header_hash, mix_hash, nonce_hex):
        if header_hash not in self.templates:
            raise ValueError("Block header hash not found in block data")          # mining.cpp:866-867
        try:
            nonce = int(nonce_hex, 16)
        except ValueError:
            raise ValueError("Invalid hex nonce") from None                          # :863-864
        t = self.templates[header_hash]
        final, mix = self.oracle.kawpow_hash(t.height // 7500, t.height, bytes.fromhex(header_hash), nonce)
        if not node.check_proof_of_work(final, t.bits):
            self.rejected.append((header_hash, nonce))
            raise ValueError("Block does not solve the boundary")                  # :880-881
        if mix.hex() != mix_hash:
            self.rejected.append((header_hash, nonce))
            return "invalid-mix-hash"                                              # validation.cpp:11660
        if t.height != self.height:
            return "inconclusive"                                                  # a solution for an old tip
        t.nonce64, t.mix_hash = nonce, mix[::-1]
        self.accepted.append(t)
        self.prev, self.height = t.block_hash(), self.height + 1
        return None                                                                # BIP22: null = accepted

    def getmininginfo(self):
        return {"blocks": self.height - 1, "chain": "mock"}

    # ---- HTTP JSON-RPC 1.0 ----
    def serve(self):
        outer = self

        class Handler(BaseHTTPRequestHandler):
            def log_message(self, *a):  # noqa: D401 - silence
                pass

            def do_POST(self):
                if self.headers.get("Authorization") != outer.auth:
                    self.send_response(401)
                    self.end_headers()
                    return
                req = json.loads(self.rfile.read(int(self.headers["Content-Length"])))
                outer.calls.append(req["method"])
                try:
                    result, error = getattr(outer, req["method"])(*req["params"]), None
                except AttributeError:
                    result, error = None, {"code": -32601, "message": "Method not found"}
                except ValueError as e:
                    result, error = None, {"code": -22, "message": str(e)}
                body = json.dumps({"result": result, "error": error, "id": req["id"]}).encode()
                self.send_response(200 if error is None else 500)
                self.send_header("Content-Type", "application/json")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

        self.server = HTTPServer(("127.0.0.1", 0), Handler)
        threading.Thread(target=self.server.serve_forever, daemon=True).start()
        return f"http://127.0.0.1:{self.server.server_address[1]}/"

    def stop(self):
        if self.server:
            self.server.shutdown()
            self.server.server_close()
