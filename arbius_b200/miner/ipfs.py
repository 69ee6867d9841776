"""Storage adapters: how a solution's files become a CID (and get pinned).

The reference uploads files to kubo or Pinata and uses the CID the service returns
(reference: miner/src/ipfs.ts:28-114). Here the CID is always computed locally by
``protocol.cid`` — identical bytes, no network round trip on the latency path — and pinning is
a best-effort side effect per strategy:
  local        compute only (optionally write blocks' files under a directory for a later `ipfs add`)
  http_client  POST /api/v0/add to a kubo daemon with the reference's options, verify the returned CID
  pinata       multipart POST pinFileToIPFS with pinataOptions {cidVersion:0}, verify IpfsHash
A mismatch between the local and the remote CID raises: it would mean the DAG layout assumptions
 are wrong for that service.
"""
from __future__ import annotations

import json
import os
import uuid
from pathlib import Path
from typing import Dict, Optional

from ..protocol.cid import cid_b58, directory_cid, file_cid
from ..protocol.config import MiningConfig

IPFS_OPTIONS = {"cid-version": "0", "hash": "sha2-256", "chunker": "size-262144", "raw-leaves": "false"}


class PinError(Exception):
    pass


def _multipart(fields: Dict[str, str], files: Dict[str, bytes], file_field: str = "file"):
    boundary = "----arbiusb200" + uuid.uuid4().hex
    body = bytearray()
    for k, v in fields.items():
        body += f"--{boundary}\r\nContent-Disposition: form-data; name=\"{k}\"\r\n\r\n{v}\r\n".encode()
    for name, data in files.items():
        body += (f"--{boundary}\r\nContent-Disposition: form-data; name=\"{file_field}\"; filename=\"{name}\"\r\n"
                 "Content-Type: application/octet-stream\r\n\r\n").encode() + data + b"\r\n"
    body += f"--{boundary}--\r\n".encode()
    return bytes(body), f"multipart/form-data; boundary={boundary}"


def _post(url: str, body: bytes, headers: Dict[str, str], timeout: float = 60.0) -> bytes:
    import urllib.request
    req = urllib.request.Request(url, data=body, headers=headers, method="POST")
    with urllib.request.urlopen(req, timeout=timeout) as r:
        return r.read()


def pin_files(c: MiningConfig, taskid: str, files: Dict[str, bytes], pin: bool = True) -> str:
    """``pinFilesToIPFS``: CIDv0 (base58) of the directory wrapping ``files`` (name -> bytes)."""
    local = cid_b58(directory_cid(files))
    strategy = c.ipfs.strategy
    if not pin or strategy == "local":
        out_dir = (c.ipfs.http_client or {}).get("spool") if strategy == "local" else None
        if out_dir:
            d = Path(out_dir) / taskid
            d.mkdir(parents=True, exist_ok=True)
            for n, b in files.items():
                (d / n).write_bytes(b)
        return local
    if strategy == "http_client":
        q = "&".join(f"{k}={v}" for k, v in IPFS_OPTIONS.items()) + "&wrap-with-directory=true&pin=true"
        body, ctype = _multipart({}, files)
        raw = _post(c.ipfs.http_client["url"].rstrip("/") + "/api/v0/add?" + q, body, {"Content-Type": ctype})
        remote = None
        for line in raw.decode().splitlines():
            ent = json.loads(line)
            if ent.get("Name") == "":
                remote = ent["Hash"]
        if remote is None:
            raise PinError("ipfs cid extract failed")
    elif strategy == "pinata":
        body, ctype = _multipart({"pinataOptions": json.dumps({"cidVersion": 0})},
                                 {f"{taskid}/{n}": b for n, b in files.items()})
        raw = _post("https://api.pinata.cloud/pinning/pinFileToIPFS", body,
                    {"Content-Type": ctype, "Authorization": f"Bearer {c.ipfs.pinata.get('jwt')}"})
        remote = json.loads(raw)["IpfsHash"]
    else:
        raise PinError("unknown ipfs strategy")
    if remote != local:
        raise PinError(f"remote CID {remote} != locally computed {local}")
    return local


def pin_file(c: MiningConfig, content: bytes, filename: str, pin: bool = True) -> str:
    """``pinFileToIPFS``: CIDv0 of a single (unwrapped) file, e.g. a task's input JSON."""
    local = cid_b58(file_cid(content))
    strategy = c.ipfs.strategy
    if not pin or strategy == "local":
        return local
    if strategy == "http_client":
        q = "&".join(f"{k}={v}" for k, v in IPFS_OPTIONS.items()) + "&pin=true"
        body, ctype = _multipart({}, {filename: content})
        raw = _post(c.ipfs.http_client["url"].rstrip("/") + "/api/v0/add?" + q, body, {"Content-Type": ctype})
        remote = json.loads(raw.decode().splitlines()[-1])["Hash"]
    elif strategy == "pinata":
        body, ctype = _multipart({"pinataOptions": json.dumps({"cidVersion": 0})}, {filename: content})
        raw = _post("https://api.pinata.cloud/pinning/pinFileToIPFS", body,
                    {"Content-Type": ctype, "Authorization": f"Bearer {c.ipfs.pinata.get('jwt')}"})
        remote = json.loads(raw)["IpfsHash"]
    else:
        raise PinError("unknown ipfs strategy")
    if remote != local:
        raise PinError(f"remote CID {remote} != locally computed {local}")
    return local
