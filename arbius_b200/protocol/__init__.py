"""Byte-level protocol contracts kept identical to the reference miner."""
from .abi import decode, decode_call, encode, encode_call, generate_commitment, hash_model, hash_task
from .base58 import b58decode, b58encode
from .cid import (cid_b58, cid_hex, directory_cid, directory_node, file_cid, file_dag, onchain_cid,
                  solution_cid, varint)
from .hydrate import (FilterResult, HydrationResult, MiningFilter, ModelEntry, check_model_filter,
                      get_model_by_id, hydrate_input)
from .keccak import event_topic, keccak256, keccak256_hex, selector
from .seed import SEED_MODULUS, taskid2seed
from .templates import TEMPLATE_NAMES, load_template, output_filename, output_mime, template_bytes, template_cid
