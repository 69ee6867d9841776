"""Model template registry. Template bytes are hashed on chain
(reference: contract/contracts/EngineV1.sol:557-575), so the files under
``templates/`` are byte-identical copies of the reference's ``templates/*.json``.
"""
from __future__ import annotations

import json
from functools import lru_cache
from pathlib import Path

from .cid import onchain_cid

TEMPLATE_DIR = Path(__file__).resolve().parents[2] / "templates"
TEMPLATE_NAMES = ("anythingv3", "kandinsky2", "damo", "zeroscopev2xl", "robust_video_matting")


@lru_cache(maxsize=None)
def template_bytes(name: str) -> bytes:
    return (TEMPLATE_DIR / f"{name}.json").read_bytes()


@lru_cache(maxsize=None)
def load_template(name: str) -> dict:
    return json.loads(template_bytes(name))


def template_cid(name: str) -> str:
    """CID the contract stores for a registered model (EngineV1.sol:564)."""
    return "0x" + onchain_cid(template_bytes(name)).hex()


def output_filename(template: dict) -> str:
    return template["output"][0]["filename"]


def output_mime(template: dict) -> str:
    fn = output_filename(template)
    kind = template["output"][0]["type"]
    ext = fn.rsplit(".", 1)[-1]
    return {"image": "image/", "video": "video/", "audio": "audio/", "text": "text/"}[kind] + ext
