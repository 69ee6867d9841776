"""Pinata garbage collection: unpin files older than a cut-off (default two hours).

The reference ships this as a fully commented-out script (reference: miner/src/scripts/pinata_unpin_old_files.ts:1-69:
``GET /data/pinList?status=pinned&pinStart=…&pinEnd=…`` then ``DELETE /pinning/unpin/<cid>`` per row). Solutions only
need to stay retrievable through the contestation window, so a miner on a metered Pinata plan wants this.

  python -m arbius_b200.miner.pinata_gc MiningConfig.json --older-than-hours 2 [--dry-run]
"""
from __future__ import annotations

import argparse
import json
import time
import urllib.request
from datetime import datetime, timezone
from typing import Callable, List, Optional

API = "https://api.pinata.cloud"


def _http(method: str, url: str, jwt: str, timeout: float = 60.0) -> bytes:
    req = urllib.request.Request(url, method=method, headers={"Authorization": f"Bearer {jwt}"})
    with urllib.request.urlopen(req, timeout=timeout) as resp:
        return resp.read()


def _iso(ts: float) -> str:
    return datetime.fromtimestamp(ts, tz=timezone.utc).strftime("%Y-%m-%dT%H:%M:%S.000Z")


def list_old_pins(jwt: str, older_than_s: float, now: Optional[float] = None, http: Callable = _http,
                  page_limit: int = 1000) -> List[str]:
    """CIDs pinned before now - older_than_s (paged)"""
    now = time.time() if now is None else now
    end = _iso(now - older_than_s)
    out, offset = [], 0
    while True:
        url = f"{API}/data/pinList?status=pinned&pinEnd={end}&pageLimit={page_limit}&pageOffset={offset}"
        rows = json.loads(http("GET", url, jwt)).get("rows", [])
        out.extend(r["ipfs_pin_hash"] for r in rows)
        if len(rows) < page_limit:
            return out
        offset += page_limit


def unpin_old_files(jwt: str, older_than_s: float = 7200.0, dry_run: bool = False, http: Callable = _http,
                    now: Optional[float] = None, log: Callable[[str], None] = print) -> int:
    cids = list_old_pins(jwt, older_than_s, now=now, http=http)
    n = 0
    for cid in cids:
        log(f"unpinning cid {cid}")
        if not dry_run:
            http("DELETE", f"{API}/pinning/unpin/{cid}", jwt)
        n += 1
    return n


def main(argv=None) -> int:
    from ..protocol.config import load_mining_config
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("config")
    ap.add_argument("--older-than-hours", type=float, default=2.0)
    ap.add_argument("--dry-run", action="store_true")
    a = ap.parse_args(argv)
    c = load_mining_config(a.config)
    jwt = c.ipfs.pinata.get("jwt")
    if not jwt:
        raise SystemExit("error: ipfs.pinata.jwt is not set in the mining config")
    n = unpin_old_files(jwt, a.older_than_hours * 3600.0, a.dry_run)
    print(f"{'would unpin' if a.dry_run else 'unpinned'} {n} files")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
