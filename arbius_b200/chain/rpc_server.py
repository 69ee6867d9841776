"""HTTP JSON-RPC front of DevChain: the eth_* subset ethers/the miner use plus Hardhat's evm_*/hardhat_* controls.

``python -m arbius_b200.chain.rpc_server --port 8545`` is this framework's ``npx hardhat node`` with the contracts
already deployed (reference: setup_local.sh:1-23; CLI tasks ``mine``/``timetravel`` contract/tasks/index.ts:36-47).
"""
from __future__ import annotations

import argparse
import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Any, Optional

from .devnode import DevChain, RpcError
from .engine_sim import ZERO_ADDR


def _hexbytes(s: Optional[str]) -> bytes:
    if not s or s == "0x":
        return b""
    return bytes.fromhex(s[2:] if s.startswith("0x") else s)


class RpcDispatcher:
    def __init__(self, chain: DevChain):
        self.chain = chain
        self.lock = threading.RLock()

    def handle(self, method: str, params: list) -> Any:
        c = self.chain
        with self.lock:
            if method == "eth_chainId":
                return hex(c.chain_id)
            if method == "net_version":
                return str(c.chain_id)
            if method == "web3_clientVersion":
                return "arbius-b200-devchain/1"
            if method == "eth_blockNumber":
                return hex(c.blocks[-1].number)
            if method in ("eth_gasPrice", "eth_maxPriorityFeePerGas"):
                return hex(10 ** 8)
            if method == "eth_feeHistory":
                return {"oldestBlock": hex(c.blocks[-1].number), "baseFeePerGas": ["0x0", "0x0"], "gasUsedRatio": [0.0], "reward": [["0x0"]]}
            if method == "eth_getBalance":
                return hex(10_000 * 10 ** 18)
            if method == "eth_accounts":
                return []
            if method == "eth_getTransactionCount":
                return hex(c.nonces.get(params[0].lower(), 0))
            if method == "eth_getCode":
                return "0x60806040" if params[0].lower() in c._sel else "0x"
            if method == "eth_estimateGas":
                tx = params[0]
                if tx.get("to"):
                    c.eth_call(tx.get("from") or ZERO_ADDR, tx["to"], _hexbytes(tx.get("data") or tx.get("input")))
                return hex(500_000)
            if method == "eth_call":
                tx = params[0]
                return "0x" + c.eth_call(tx.get("from") or ZERO_ADDR, tx["to"], _hexbytes(tx.get("data") or tx.get("input"))).hex()
            if method == "eth_sendRawTransaction":
                return c.send_raw_transaction(_hexbytes(params[0]))
            if method == "eth_getTransactionByHash":
                return c.txs.get(params[0].lower())
            if method == "eth_getTransactionReceipt":
                return c.receipts.get(params[0].lower())
            if method == "eth_getBlockByNumber":
                b = c.block_by_number(params[0])
                return c.block_json(b, bool(params[1]) if len(params) > 1 else False) if b else None
            if method == "eth_getLogs":
                return c.get_logs(params[0])
            if method == "evm_increaseTime":
                c.increase_time(int(params[0], 16) if isinstance(params[0], str) else int(params[0]))
                return hex(c.engine.now)
            if method == "evm_setNextBlockTimestamp":
                ts = int(params[0], 16) if isinstance(params[0], str) else int(params[0])
                c.engine.now = max(c.engine.now, ts - 1)
                return None
            if method == "evm_mine":
                c.mine(1)
                return "0x0"
            if method == "hardhat_mine":
                n = int(params[0], 16) if params else 1
                interval = int(params[1], 16) if len(params) > 1 else 1
                c.mine(n, interval)
                return None
            if method == "evm_snapshot":
                return hex(c.snapshot())
            if method == "evm_revert":
                return c.revert(int(params[0], 16))
            if method == "arbius_getAddresses":
                return c.addresses
            if method == "arbius_mint":                       # faucet: devchain only
                c.token.mint(params[0], int(params[1], 16) if isinstance(params[1], str) else int(params[1]))
                return True
        raise RpcError(-32601, f"the method {method} does not exist/is not available")

    def handle_json(self, req: Any) -> Any:
        if isinstance(req, list):
            return [self.handle_json(r) for r in req]
        rid = req.get("id")
        try:
            return {"jsonrpc": "2.0", "id": rid, "result": self.handle(req.get("method", ""), req.get("params") or [])}
        except RpcError as e:
            err = {"code": e.code, "message": e.message}
            if e.data is not None:
                err["data"] = e.data
            return {"jsonrpc": "2.0", "id": rid, "error": err}
        except Exception as e:                                 # noqa: BLE001 - a node must answer, not die
            return {"jsonrpc": "2.0", "id": rid, "error": {"code": -32603, "message": f"internal error: {e!r}"}}


def make_server(chain: DevChain, host: str = "127.0.0.1", port: int = 8545) -> ThreadingHTTPServer:
    disp = RpcDispatcher(chain)

    class Handler(BaseHTTPRequestHandler):
        def do_POST(self):                                     # noqa: N802
            n = int(self.headers.get("Content-Length", "0"))
            try:
                req = json.loads(self.rfile.read(n) or b"{}")
                body = json.dumps(disp.handle_json(req)).encode()
            except json.JSONDecodeError:
                body = json.dumps({"jsonrpc": "2.0", "id": None, "error": {"code": -32700, "message": "parse error"}}).encode()
            self.send_response(200)
            self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

        def log_message(self, *a):                             # quiet
            pass

    srv = ThreadingHTTPServer((host, port), Handler)
    srv.dispatcher = disp
    return srv


def serve_in_thread(chain: DevChain, host: str = "127.0.0.1", port: int = 0):
    """-> (server, url); port 0 picks a free port (tests)"""
    srv = make_server(chain, host, port)
    th = threading.Thread(target=srv.serve_forever, daemon=True)
    th.start()
    return srv, f"http://{host}:{srv.server_address[1]}"


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="Arbius devchain: EngineV1/BaseToken/Governor behind eth JSON-RPC")
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=8545)
    ap.add_argument("--chain-id", type=int, default=31337)
    ap.add_argument("--deployer", default="0xf39fd6e51aad88f6f4ce6ab8827279cfffb92266", help="owner of engine and token")
    ap.add_argument("--fund", action="append", default=[], metavar="ADDR=AIUS", help="mint whole tokens to an address")
    a = ap.parse_args(argv)
    funded = {}
    for item in a.fund:
        addr, amt = item.split("=")
        funded[addr.lower()] = int(float(amt) * 10 ** 18)
    chain = DevChain(chain_id=a.chain_id, deployer=a.deployer, funded=funded)
    srv = make_server(chain, a.host, a.port)
    print(json.dumps({"listening": f"http://{a.host}:{a.port}", **chain.addresses}), flush=True)
    try:
        srv.serve_forever()
    except KeyboardInterrupt:
        pass
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
