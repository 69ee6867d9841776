"""IPFS CIDv0 construction (dag-pb + UnixFS) without a daemon.

The reference delegates hashing to kubo/Pinata with
``{cidVersion:0, hashAlg:'sha2-256', chunker:'size-262144', rawLeaves:false}``
and ``wrapWithDirectory:true`` (reference: miner/src/ipfs.ts:11-16, 30-50), and
re-implements the single-block case on chain
(contract/contracts/libraries/IPFS.sol:38-65). This module builds the same
bytes locally so a solution CID never needs a network round trip:

* ``onchain_cid``      - exact restatement of IPFS.sol (<= 65536 B)
* ``file_dag``         - kubo balanced layout: 262144-B leaves, 174 links / node
* ``directory_cid``    - the wrapping directory the miner submits

An optional native SHA-256/DAG builder (csrc/host/cid_native.cpp) is used for
large files when it has been built; results are identical.
"""
from __future__ import annotations

import hashlib
from dataclasses import dataclass
from typing import Iterable, Mapping

from .base58 import b58encode

CHUNK_SIZE = 262144
MAX_LINKS = 174
ONCHAIN_MAX = 65536


def varint(n: int) -> bytes:
    if n < 0:
        raise ValueError("varint of negative")
    out = bytearray()
    while n > 0x7F:
        out.append(0x80 | (n & 0x7F))
        n >>= 7
    out.append(n)
    return bytes(out)


def _mh(block: bytes) -> bytes:
    return b"\x12\x20" + hashlib.sha256(block).digest()


def onchain_cid(content: bytes) -> bytes:
    """``getIPFSCID`` of IPFS.sol:38-65 — 34 bytes ``0x1220 || sha256(...)``."""
    if len(content) > ONCHAIN_MAX:
        raise ValueError("Max content size is 65536 bytes")
    n = varint(len(content))
    meat = b"\x08\x02\x12" + n + content + b"\x18" + n
    return _mh(b"\x0a" + varint(len(meat)) + meat)


@dataclass(frozen=True)
class DagNode:
    multihash: bytes   # 34 bytes
    tsize: int         # cumulative serialized size of the sub-DAG
    filesize: int      # payload bytes covered
    block_len: int


def _leaf(chunk: bytes) -> DagNode:
    n = len(chunk)
    if n == 0:
        unixfs = b"\x08\x02\x18\x00"          # kubo omits the empty Data field
    else:
        v = varint(n)
        unixfs = b"\x08\x02\x12" + v + chunk + b"\x18" + v
    block = b"\x0a" + varint(len(unixfs)) + unixfs
    return DagNode(_mh(block), len(block), n, len(block))


def _link(child: DagNode, name: bytes) -> bytes:
    body = b"\x0a\x22" + child.multihash + b"\x12" + varint(len(name)) + name + b"\x18" + varint(child.tsize)
    return b"\x12" + varint(len(body)) + body


def _parent(children: list[DagNode]) -> DagNode:
    total = sum(c.filesize for c in children)
    unixfs = b"\x08\x02\x18" + varint(total) + b"".join(b"\x20" + varint(c.filesize) for c in children)
    block = b"".join(_link(c, b"") for c in children) + b"\x0a" + varint(len(unixfs)) + unixfs
    return DagNode(_mh(block), len(block) + sum(c.tsize for c in children), total, len(block))


def file_dag(content: bytes, chunk_size: int = CHUNK_SIZE, max_links: int = MAX_LINKS) -> DagNode:
    """Root node of the balanced UnixFS DAG kubo builds for ``content``."""
    if len(content) <= chunk_size:
        return _leaf(content)
    level = [_leaf(content[i:i + chunk_size]) for i in range(0, len(content), chunk_size)]
    while len(level) > 1:
        level = [_parent(level[i:i + max_links]) for i in range(0, len(level), max_links)]
    return level[0]


def file_cid(content: bytes) -> bytes:
    return file_dag(content).multihash


def directory_node(files: Mapping[str, bytes] | Iterable[tuple[str, bytes]]) -> DagNode:
    """Wrapping directory with one link per file, sorted by name."""
    items = dict(files)
    links = []
    tsum = 0
    for name in sorted(items, key=lambda s: s.encode()):
        root = file_dag(items[name])
        links.append(_link(root, name.encode()))
        tsum += root.tsize
    block = b"".join(links) + b"\x0a\x02\x08\x01"
    return DagNode(_mh(block), len(block) + tsum, 0, len(block))


def directory_cid(files: Mapping[str, bytes] | Iterable[tuple[str, bytes]]) -> bytes:
    return directory_node(files).multihash


def cid_hex(mh: bytes) -> str:
    """On-chain presentation: ``0x1220…`` (reference: miner/src/models.ts:52)."""
    return "0x" + mh.hex()


def cid_b58(mh: bytes) -> str:
    """User-facing ``Qm…`` form (reference: website/src/utils.ts:22-27)."""
    return b58encode(mh)


def solution_cid(files: Mapping[str, bytes]) -> str:
    """What ``default__getcid`` returns for a set of output files (models.ts:34-54)."""
    return cid_hex(directory_cid(files))
