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


def split_types(inner: str) -> list[str]:
    """split a comma list at nesting depth 0: ``"uint256,(address,bytes)[],bool"`` -> 3 entries"""
    out, depth, cur = [], 0, ""
    for ch in inner:
        if ch == "," and depth == 0:
            out.append(cur)
            cur = ""
            continue
        depth += ch == "("
        depth -= ch == ")"
        cur += ch
    if cur:
        out.append(cur)
    return out


def _array_parts(t: str):
    """``T[]`` -> (T, None); ``T[3]`` -> (T, 3); otherwise None"""
    if not t.endswith("]"):
        return None
    i = t.rindex("[")
    n = t[i + 1:-1]
    return t[:i], (int(n) if n else None)


def _is_dynamic(t: str) -> bool:
    if t in ("bytes", "string"):
        return True
    arr = _array_parts(t)
    if arr is not None:
        return arr[1] is None or _is_dynamic(arr[0])
    if t.startswith("("):
        return any(_is_dynamic(x) for x in split_types(t[1:-1]))
    return False


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


def _enc_value(t: str, v: Any) -> bytes:
    """encoding of one value (head-less): static types give their words, dynamic ones their tail blob"""
    if t in ("bytes", "string"):
        b = v.encode() if (t == "string" and isinstance(v, str)) else _to_bytes(v)
        return len(b).to_bytes(32, "big") + b + b"\x00" * ((-len(b)) % 32)
    arr = _array_parts(t)
    if arr is not None:
        inner, n = arr
        if n is not None and len(v) != n:
            raise ValueError(f"{t} needs {n} items")
        body = encode([inner] * len(v), list(v))
        return (len(v).to_bytes(32, "big") if n is None else b"") + body
    if t.startswith("("):
        return encode(split_types(t[1:-1]), list(v))
    return _enc_static(t, v)


def encode(types: Sequence[str], values: Sequence[Any]) -> bytes:
    """``abi.encode(...)``: static words, bytes/string, arrays (``T[]``, ``T[k]``) and tuples (``(a,b)``)."""
    if len(types) != len(values):
        raise ValueError("types/values length mismatch")
    blobs = [_enc_value(t, v) for t, v in zip(types, values)]
    head_size = sum(32 if _is_dynamic(t) else len(b) for t, b in zip(types, blobs))
    head, tail = b"", b""
    for t, b in zip(types, blobs):
        if _is_dynamic(t):
            head += (head_size + len(tail)).to_bytes(32, "big")
            tail += b
        else:
            head += b
    return head + tail


def _static_size(t: str) -> int:
    arr = _array_parts(t)
    if arr is not None:
        return arr[1] * _static_size(arr[0])
    if t.startswith("("):
        return sum(_static_size(x) for x in split_types(t[1:-1]))
    return 32


def _dec_value(t: str, data: bytes, pos: int) -> Any:
    """decode the value of type ``t`` whose encoding starts at ``pos``"""
    if t in ("bytes", "string"):
        n = int.from_bytes(data[pos:pos + 32], "big")
        raw = data[pos + 32:pos + 32 + n]
        if len(raw) != n:
            raise ValueError("ABI dynamic data truncated")
        return raw.decode() if t == "string" else raw
    arr = _array_parts(t)
    if arr is not None:
        inner, n = arr
        if n is None:
            n = int.from_bytes(data[pos:pos + 32], "big")
            pos += 32
        if n > len(data):
            raise ValueError("ABI array length exceeds the data")
        return _dec_tuple([inner] * n, data, pos)
    if t.startswith("("):
        return _dec_tuple(split_types(t[1:-1]), data, pos)
    word = data[pos:pos + 32]
    if len(word) != 32:
        raise ValueError("ABI data too short")
    if t == "address":
        return "0x" + word[12:].hex()
    if t == "bool":
        return bool(int.from_bytes(word, "big"))
    if t.startswith("uint"):
        return int.from_bytes(word, "big")
    if t.startswith("int"):
        iv = int.from_bytes(word, "big")
        return iv - (1 << 256) if iv >> 255 else iv
    if t.startswith("bytes"):
        return word[:int(t[5:])]
    raise ValueError(f"unsupported ABI type {t}")


def _dec_tuple(types: Sequence[str], data: bytes, base: int) -> list[Any]:
    out, pos = [], base
    for t in types:
        if _is_dynamic(t):
            word = data[pos:pos + 32]
            if len(word) != 32:
                raise ValueError("ABI data too short")
            out.append(_dec_value(t, data, base + int.from_bytes(word, "big")))
            pos += 32
        else:
            out.append(_dec_value(t, data, pos))
            pos += _static_size(t)
    return out


def decode(types: Sequence[str], data: bytes) -> list[Any]:
    return _dec_tuple(list(types), bytes(data), 0)


def encode_call(signature: str, values: Sequence[Any]) -> bytes:
    """Selector + encoded args for e.g. ``submitTask(uint8,address,bytes32,uint256,bytes)``."""
    inner = signature[signature.index("(") + 1:signature.rindex(")")]
    return selector(signature) + encode(split_types(inner), values)


def decode_call(signature: str, calldata: bytes) -> list[Any]:
    if calldata[:4] != selector(signature):
        raise ValueError("selector mismatch")
    inner = signature[signature.index("(") + 1:signature.rindex(")")]
    return decode(split_types(inner), calldata[4:])


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
