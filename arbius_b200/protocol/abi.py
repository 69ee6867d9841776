"""Minimal Solidity ABI encoder/decoder (the subset the Engine flow needs).

Covers the head/tail layout for static words (address, uintN, intN, bytes32, bool)
and dynamic ``bytes``/``string``. That is what ``generateCommitment``
(reference: miner/src/utils.ts:42-49), ``hashTask``/``hashModel``
(contract/contracts/EngineV1.sol:419-438), and the submitTask/submitSolution
call data use (miner/src/index.ts:151-159, 483-501, 632-640).
"""
from __future__ import annotations

from typing import Any, Sequence

from .keccak import keccak256, selector


def _to_bytes(v: Any) -> bytes:
    if isinstance(v, (bytes, bytearray)):
        return bytes(v)
    if isinstance(v, str):
        s = v[2:] if v.startswith(("0x", "0X")) else v
        if len(s) % 2:
            s = "0" + s
        return bytes.fromhex(s)
    raise TypeError(f"cannot convert {type(v)} to bytes")


def _is_dynamic(t: str) -> bool:
    return t in ("bytes", "string")


def _enc_static(t: str, v: Any) -> bytes:
    if t == "address":
        b = _to_bytes(v)
        if len(b) != 20:
            raise ValueError("address must be 20 bytes")
        return b.rjust(32, b"\x00")
    if t == "bool":
        return int(bool(v)).to_bytes(32, "big")
    if t.startswith("uint"):
        iv = int(v, 0) if isinstance(v, str) else int(v)
        bits = int(t[4:] or 256)
        if iv < 0 or iv >= 1 << bits:
            raise ValueError(f"{t} out of range")
        return iv.to_bytes(32, "big")
    if t.startswith("int"):
        iv = int(v, 0) if isinstance(v, str) else int(v)
        return (iv % (1 << 256)).to_bytes(32, "big")
    if t.startswith("bytes"):
        n = int(t[5:])
        b = _to_bytes(v)
        if len(b) != n:
            raise ValueError(f"{t} needs {n} bytes, got {len(b)}")
        return b.ljust(32, b"\x00")
    raise ValueError(f"unsupported ABI type {t}")


def _enc_dynamic(t: str, v: Any) -> bytes:
    b = v.encode() if (t == "string" and isinstance(v, str)) else _to_bytes(v)
    pad = (-len(b)) % 32
    return len(b).to_bytes(32, "big") + b + b"\x00" * pad


def encode(types: Sequence[str], values: Sequence[Any]) -> bytes:
    """``abi.encode(...)`` for a flat tuple of supported types."""
    if len(types) != len(values):
        raise ValueError("types/values length mismatch")
    head_size = 32 * len(types)
    head, tail = b"", b""
    for t, v in zip(types, values):
        if _is_dynamic(t):
            head += (head_size + len(tail)).to_bytes(32, "big")
            tail += _enc_dynamic(t, v)
        else:
            head += _enc_static(t, v)
    return head + tail


def decode(types: Sequence[str], data: bytes) -> list[Any]:
    out: list[Any] = []
    for i, t in enumerate(types):
        word = data[32 * i:32 * i + 32]
        if len(word) != 32:
            raise ValueError("ABI data too short")
        if _is_dynamic(t):
            off = int.from_bytes(word, "big")
            n = int.from_bytes(data[off:off + 32], "big")
            raw = data[off + 32:off + 32 + n]
            if len(raw) != n:
                raise ValueError("ABI dynamic data truncated")
            out.append(raw.decode() if t == "string" else raw)
        elif t == "address":
            out.append("0x" + word[12:].hex())
        elif t == "bool":
            out.append(bool(int.from_bytes(word, "big")))
        elif t.startswith("uint"):
            out.append(int.from_bytes(word, "big"))
        elif t.startswith("int"):
            iv = int.from_bytes(word, "big")
            out.append(iv - (1 << 256) if iv >> 255 else iv)
        elif t.startswith("bytes"):
            out.append(word[:int(t[5:])])
        else:
            raise ValueError(f"unsupported ABI type {t}")
    return out


def encode_call(signature: str, values: Sequence[Any]) -> bytes:
    """Selector + encoded args for e.g. ``submitTask(uint8,address,bytes32,uint256,bytes)``."""
    inner = signature[signature.index("(") + 1:signature.rindex(")")]
    types = [t for t in inner.split(",") if t]
    return selector(signature) + encode(types, values)


def decode_call(signature: str, calldata: bytes) -> list[Any]:
    if calldata[:4] != selector(signature):
        raise ValueError("selector mismatch")
    inner = signature[signature.index("(") + 1:signature.rindex(")")]
    types = [t for t in inner.split(",") if t]
    return decode(types, calldata[4:])


def generate_commitment(address: Any, taskid: Any, cid: Any) -> str:
    """keccak256(abi.encode(address, bytes32 taskid, bytes cid)) as 0x-hex.

    Reference: miner/src/utils.ts:42-49 == EngineV1.sol:537-543.
    """
    return "0x" + keccak256(encode(["address", "bytes32", "bytes"], [address, taskid, cid])).hex()


def hash_model(addr: Any, fee: int, rate: int, cid: Any, sender: Any) -> str:
    """EngineV1.hashModel: keccak256(abi.encode(sender, addr, fee, cid)) (EngineV1.sol:419-424).

    ``rate`` is part of the struct but NOT of the hash.
    """
    del rate
    return "0x" + keccak256(encode(["address", "address", "uint256", "bytes"], [sender, addr, fee, cid])).hex()


def hash_task(model: Any, fee: int, cid: Any, sender: Any, prevhash: Any) -> str:
    """EngineV1.hashTask: keccak256(abi.encode(sender, prevhash, model, fee, cid)) (EngineV1.sol:431-438)."""
    return "0x" + keccak256(
        encode(["address", "bytes32", "bytes32", "uint256", "bytes"], [sender, prevhash, model, fee, cid])
    ).hex()
