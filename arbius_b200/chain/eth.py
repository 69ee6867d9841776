"""Ethereum wire primitives in pure Python: secp256k1 (RFC 6979 signing, public-key recovery), RLP, and legacy
(EIP-155) / EIP-1559 transaction encoding. This is what ethers' ``Wallet`` does for the reference miner
(reference: miner/src/blockchain.ts:22-30); no native dependency, integers only.
"""
from __future__ import annotations

import hashlib
import hmac
from dataclasses import dataclass
from typing import Any, List, Optional, Tuple, Union

from ..protocol.keccak import keccak256

# ------------------------------------------------------------------------------------------- secp256k1
P = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEFFFFFC2F
N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141
GX = 0x79BE667EF9DCBBAC55A06295CE870B07029BFCDB2DCE28D959F2815B16F81798
GY = 0x483ADA7726A3C4655DA4FBFC0E1108A8FD17B448A68554199C47D08FFB10D4B8
_INF = None


def _add(a, b):
    if a is _INF:
        return b
    if b is _INF:
        return a
    x1, y1 = a
    x2, y2 = b
    if x1 == x2:
        if (y1 + y2) % P == 0:
            return _INF
        lam = 3 * x1 * x1 * pow(2 * y1, -1, P) % P
    else:
        lam = (y2 - y1) * pow(x2 - x1, -1, P) % P
    x3 = (lam * lam - x1 - x2) % P
    return x3, (lam * (x1 - x3) - y1) % P


def _mul(k: int, pt):
    out = _INF
    while k:
        if k & 1:
            out = _add(out, pt)
        pt = _add(pt, pt)
        k >>= 1
    return out


def _priv_int(priv: Union[str, bytes, int]) -> int:
    if isinstance(priv, int):
        d = priv
    elif isinstance(priv, bytes):
        d = int.from_bytes(priv, "big")
    else:
        d = int(priv[2:] if priv.startswith("0x") else priv, 16)
    if not 0 < d < N:
        raise ValueError("private key out of range")
    return d


def public_key(priv) -> bytes:
    """64-byte uncompressed public key (x ‖ y)"""
    x, y = _mul(_priv_int(priv), (GX, GY))
    return x.to_bytes(32, "big") + y.to_bytes(32, "big")


def address_of_pubkey(pub: bytes) -> str:
    return "0x" + keccak256(pub)[12:].hex()


def address_of(priv) -> str:
    return address_of_pubkey(public_key(priv))


def to_checksum_address(addr: str) -> str:
    a = addr.lower().replace("0x", "")
    h = keccak256(a.encode()).hex()
    return "0x" + "".join(c.upper() if int(h[i], 16) >= 8 else c for i, c in enumerate(a))


def _rfc6979(d: int, h: bytes) -> int:
    x = d.to_bytes(32, "big")
    v, k = b"\x01" * 32, b"\x00" * 32
    k = hmac.new(k, v + b"\x00" + x + h, hashlib.sha256).digest()
    v = hmac.new(k, v, hashlib.sha256).digest()
    k = hmac.new(k, v + b"\x01" + x + h, hashlib.sha256).digest()
    v = hmac.new(k, v, hashlib.sha256).digest()
    while True:
        v = hmac.new(k, v, hashlib.sha256).digest()
        cand = int.from_bytes(v, "big")
        if 0 < cand < N:
            return cand
        k = hmac.new(k, v + b"\x00", hashlib.sha256).digest()
        v = hmac.new(k, v, hashlib.sha256).digest()


def sign(priv, msg_hash: bytes) -> Tuple[int, int, int]:
    """(recovery id 0/1, r, s) with low-s normalisation (EIP-2)"""
    d = _priv_int(priv)
    z = int.from_bytes(msg_hash, "big")
    k = _rfc6979(d, msg_hash)
    x, y = _mul(k, (GX, GY))
    r = x % N
    s = pow(k, -1, N) * (z + r * d) % N
    if r == 0 or s == 0:                                         # pragma: no cover - probability 2^-256
        raise ValueError("degenerate signature")
    rec = (y & 1) | (2 if x >= N else 0)
    if s > N // 2:
        s = N - s
        rec ^= 1
    return rec, r, s


def recover(msg_hash: bytes, rec: int, r: int, s: int) -> bytes:
    """64-byte public key of the signer"""
    if not (0 < r < N and 0 < s < N and rec in (0, 1, 2, 3)):
        raise ValueError("bad signature")
    x = r + (N if rec & 2 else 0)
    if x >= P:
        raise ValueError("bad signature")
    y2 = (pow(x, 3, P) + 7) % P
    y = pow(y2, (P + 1) // 4, P)
    if y * y % P != y2:
        raise ValueError("bad signature: not on curve")
    if (y & 1) != (rec & 1):
        y = P - y
    z = int.from_bytes(msg_hash, "big")
    rinv = pow(r, -1, N)
    q = _add(_mul(s * rinv % N, (x, y)), _mul((-z * rinv) % N, (GX, GY)))
    if q is _INF:
        raise ValueError("bad signature")
    return q[0].to_bytes(32, "big") + q[1].to_bytes(32, "big")


# ------------------------------------------------------------------------------------------------ RLP
def _int_bytes(v: int) -> bytes:
    return b"" if v == 0 else v.to_bytes((v.bit_length() + 7) // 8, "big")


def _to_item(v: Any) -> Union[bytes, list]:
    if isinstance(v, (list, tuple)):
        return [_to_item(x) for x in v]
    if isinstance(v, bool):
        return b"\x01" if v else b""
    if isinstance(v, int):
        return _int_bytes(v)
    if isinstance(v, str):
        h = v[2:] if v.startswith("0x") else v
        return bytes.fromhex(h if len(h) % 2 == 0 else "0" + h)
    return bytes(v)


def _len_prefix(n: int, base: int) -> bytes:
    if n < 56:
        return bytes([base + n])
    lb = _int_bytes(n)
    return bytes([base + 55 + len(lb)]) + lb


def rlp_encode(v: Any) -> bytes:
    it = _to_item(v)
    if isinstance(it, list):
        body = b"".join(rlp_encode(x) for x in it)
        return _len_prefix(len(body), 0xC0) + body
    if len(it) == 1 and it[0] < 0x80:
        return it
    return _len_prefix(len(it), 0x80) + it


def _decode_at(data: bytes, pos: int):
    b0 = data[pos]
    if b0 < 0x80:
        return data[pos:pos + 1], pos + 1
    if b0 < 0xB8:
        n = b0 - 0x80
        return data[pos + 1:pos + 1 + n], pos + 1 + n
    if b0 < 0xC0:
        ll = b0 - 0xB7
        n = int.from_bytes(data[pos + 1:pos + 1 + ll], "big")
        return data[pos + 1 + ll:pos + 1 + ll + n], pos + 1 + ll + n
    if b0 < 0xF8:
        n, start = b0 - 0xC0, pos + 1
    else:
        ll = b0 - 0xF7
        n, start = int.from_bytes(data[pos + 1:pos + 1 + ll], "big"), pos + 1 + ll
    end, out, p = start + n, [], start
    while p < end:
        item, p = _decode_at(data, p)
        out.append(item)
    if p != end:
        raise ValueError("RLP list overruns its length")
    return out, end


def rlp_decode(data: bytes):
    item, end = _decode_at(data, 0)
    if end != len(data):
        raise ValueError("trailing bytes after RLP item")
    return item


# ------------------------------------------------------------------------------------------ transactions
@dataclass
class Transaction:
    chain_id: int
    nonce: int
    to: Optional[str]
    value: int = 0
    data: bytes = b""
    gas: int = 21000
    gas_price: Optional[int] = None                    # legacy
    max_fee_per_gas: Optional[int] = None              # EIP-1559
    max_priority_fee_per_gas: Optional[int] = None
    sender: Optional[str] = None                       # filled by decode_raw_transaction
    hash: Optional[str] = None

    @property
    def is_1559(self) -> bool:
        return self.max_fee_per_gas is not None

    def _fields(self) -> list:
        to = bytes.fromhex(self.to[2:]) if self.to else b""
        if self.is_1559:
            return [self.chain_id, self.nonce, self.max_priority_fee_per_gas or 0, self.max_fee_per_gas, self.gas, to,
                    self.value, self.data, []]
        return [self.nonce, self.gas_price or 0, self.gas, to, self.value, self.data]

    def signing_hash(self) -> bytes:
        if self.is_1559:
            return keccak256(b"\x02" + rlp_encode(self._fields()))
        return keccak256(rlp_encode(self._fields() + [self.chain_id, 0, 0]))


def sign_transaction(tx: Transaction, priv) -> bytes:
    """raw signed transaction bytes (what eth_sendRawTransaction takes)"""
    rec, r, s = sign(priv, tx.signing_hash())
    if tx.is_1559:
        return b"\x02" + rlp_encode(tx._fields() + [rec, r, s])
    return rlp_encode(tx._fields() + [tx.chain_id * 2 + 35 + rec, r, s])


def _i(b: bytes) -> int:
    return int.from_bytes(b, "big")


def decode_raw_transaction(raw: bytes) -> Transaction:
    """parse + recover the sender of a signed legacy/EIP-155 or type-2 transaction"""
    if raw[:1] == b"\x02":
        f = rlp_decode(raw[1:])
        if len(f) != 12:
            raise ValueError("malformed type-2 transaction")
        tx = Transaction(chain_id=_i(f[0]), nonce=_i(f[1]), max_priority_fee_per_gas=_i(f[2]), max_fee_per_gas=_i(f[3]),
                         gas=_i(f[4]), to=("0x" + f[5].hex()) if f[5] else None, value=_i(f[6]), data=f[7])
        rec, r, s = _i(f[9]), _i(f[10]), _i(f[11])
    else:
        f = rlp_decode(raw)
        if not isinstance(f, list) or len(f) != 9:
            raise ValueError("malformed legacy transaction")
        v, r, s = _i(f[6]), _i(f[7]), _i(f[8])
        if v in (27, 28):
            raise ValueError("pre-EIP-155 transactions are not accepted")
        chain_id = (v - 35) // 2
        rec = (v - 35) % 2
        tx = Transaction(chain_id=chain_id, nonce=_i(f[0]), gas_price=_i(f[1]), gas=_i(f[2]),
                         to=("0x" + f[3].hex()) if f[3] else None, value=_i(f[4]), data=f[5])
    tx.sender = address_of_pubkey(recover(tx.signing_hash(), rec, r, s))
    tx.hash = "0x" + keccak256(raw).hex()
    return tx


def generate_wallet(entropy: Optional[bytes] = None) -> Tuple[str, str]:
    """(private key hex, address) — the ``gen-wallet`` task (reference: contract/tasks/index.ts:49-54)"""
    import os
    while True:
        raw = hashlib.sha256(entropy).digest() if entropy is not None else os.urandom(32)
        d = int.from_bytes(raw, "big")
        if 0 < d < N:
            return "0x" + raw.hex(), address_of(d)
        entropy = raw                                            # pragma: no cover
