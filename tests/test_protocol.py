"""Golden-vector tests for the byte-level contracts (SURVEY §4 carry-overs)."""
import hashlib
import json
from pathlib import Path

import pytest
from hypothesis import given, settings, strategies as st

from arbius_b200.protocol import (b58decode, b58encode, check_model_filter, cid_b58, cid_hex, decode,
                                  decode_call, directory_cid, directory_node, encode, encode_call,
                                  file_cid, file_dag, generate_commitment, hash_task, hydrate_input,
                                  keccak256_hex, load_template, onchain_cid, selector, solution_cid,
                                  taskid2seed, template_cid, varint, MiningFilter, ModelEntry,
                                  TEMPLATE_NAMES, template_bytes)
from arbius_b200.protocol.config import load_mining_config, STATIC_CHAIN_CONFIG

DATA = Path(__file__).parent / "data"

# contract/test/ipfs.ts:52-55 == miner/test/ipfs.test.ts:87-135
IPFS_GOLDEN = {
    "a": "0x1220e844b8764c00d4a76ac03930a3d8f32f3df59aea3ed0ade4c3bc38a3b23a31d9",
    "b": "0x1220f782bf27d7dfa16c5556ae0e19d41a73fc380a28455abcedecd70460505f022b",
    "c": "0x1220c32cae42b7d6ed6efd2512fd7dac6530cbd96cbcc19a3d1c336ace8e401f1c3a",
    "d": "0x1220f4ad8a3bd3189da2ad909ee41148d6893d8c629c410f7f2c7e3fae75aade79c8",
}


@pytest.mark.parametrize("name", list(IPFS_GOLDEN))
def test_ipfs_goldens(name):
    data = (DATA / f"ipfs_{name}.bin").read_bytes()
    assert cid_hex(onchain_cid(data)) == IPFS_GOLDEN[name]
    assert cid_hex(file_cid(data)) == IPFS_GOLDEN[name]


def test_testbuf_cid():
    # contract/test/base.test.ts:9-10 TESTBUF="test\n"
    assert cid_hex(onchain_cid(bytes.fromhex("746573740a"))) == IPFS_GOLDEN["d"]


def test_kandinsky_template_cid_matches_chain_config():
    assert template_cid("kandinsky2") == STATIC_CHAIN_CONFIG["models"]["kandinsky2"]["params"]["cid"]


def test_templates_are_byte_identical_to_known_sizes():
    sizes = {"anythingv3": 1854, "kandinsky2": 760, "damo": 880, "zeroscopev2xl": 1446,
             "robust_video_matting": 793}
    for n in TEMPLATE_NAMES:
        assert len(template_bytes(n)) == sizes[n]
        t = load_template(n)
        assert t["output"][0]["filename"].startswith("out-1.")


def test_known_kubo_cids():
    assert cid_b58(file_cid(b"")) == "QmbFMke1KXqnYyBBWxB74N4c5SBnJMVAiMNRcGu6x1AwQH"
    assert cid_b58(file_cid(b"hello world\n")) == "QmT78zSuBmuS4z925WZfrqQ1qHaJ56DQaTfyMUF7F8ff5o"
    assert cid_b58(directory_cid({})) == "QmUNLLsPACCz1vLxQVkXqqLX5R1X345qqfHbsf67hvA3Nn"


def test_onchain_limit():
    with pytest.raises(ValueError):
        onchain_cid(b"\0" * 65537)


def test_multichunk_structure():
    data = bytes(range(256)) * 4096  # 1 MiB -> 4 leaves
    root = file_dag(data)
    assert root.filesize == len(data)
    leaves = [file_dag(data[i:i + 262144]) for i in range(0, len(data), 262144)]
    assert len(leaves) == 4
    # rebuild parent by hand
    links = b""
    for lf in leaves:
        body = b"\x0a\x22" + lf.multihash + b"\x12\x00" + b"\x18" + varint(lf.tsize)
        links += b"\x12" + varint(len(body)) + body
    unixfs = b"\x08\x02\x18" + varint(len(data)) + b"".join(b"\x20" + varint(262144) for _ in range(4))
    block = links + b"\x0a" + varint(len(unixfs)) + unixfs
    assert root.multihash == b"\x12\x20" + hashlib.sha256(block).digest()
    assert root.tsize == len(block) + sum(l.tsize for l in leaves)


def test_deep_tree_over_174_links():
    data = b"\x01" * (262144 * 176 + 5)
    root = file_dag(data, chunk_size=262144)
    assert root.filesize == len(data)
    small = file_dag(b"ab" * 700, chunk_size=4, max_links=3)  # forces 3+ levels
    assert small.filesize == 1400


def test_directory_wrap_and_solution_cid():
    png = b"\x89PNG" + b"x" * 1000
    node = directory_node({"out-1.png": png})
    f = file_dag(png)
    body = b"\x0a\x22" + f.multihash + b"\x12\x09out-1.png\x18" + varint(f.tsize)
    block = b"\x12" + varint(len(body)) + body + b"\x0a\x02\x08\x01"
    assert node.multihash == b"\x12\x20" + hashlib.sha256(block).digest()
    s = solution_cid({"out-1.png": png})
    assert s.startswith("0x1220") and len(s) == 70
    assert solution_cid({"b": b"1", "a": b"2"}) == solution_cid({"a": b"2", "b": b"1"})


@given(st.binary(max_size=200))
@settings(max_examples=50, deadline=None)
def test_base58_roundtrip(b):
    assert b58decode(b58encode(b)) == b


def test_base58_cid_form():
    mh = bytes.fromhex(IPFS_GOLDEN["d"][2:])
    s = b58encode(mh)
    assert s.startswith("Qm") and len(s) == 46 and b58decode(s) == mh


def test_keccak_vectors():
    assert keccak256_hex(b"") == "0xc5d2460186f7233c927e7db2dcc703c0e500b653ca82273b7bfad8045d85a470"
    assert keccak256_hex(b"abc") == "0x4e03657aea45a94fc7d47ba826c8d667c0d1e6e33a64a036ec44f58fa12d6c45"
    assert keccak256_hex(b"a" * 200) == keccak256_hex(b"a" * 200)
    assert selector("transfer(address,uint256)").hex() == "a9059cbb"
    assert selector("balanceOf(address)").hex() == "70a08231"


def test_commitment_encoding_layout():
    # triple from miner/test/utils.test.ts:61-63
    address = "0x1A320E53A25f518B893F286f3600cc204c181a8E"
    taskid = "0xdc6a147f2cd937a1b290b0dc4eff49a084ad72468fcb36df6d8ddb00c5ff6f7b"
    cid = "0x122002e2550d45270ed9c0df80be9a331940d391bcb138e0edfce4d2ff20168d6691"
    enc = encode(["address", "bytes32", "bytes"], [address, taskid, cid])
    assert len(enc) == 32 * 3 + 32 + 64
    assert enc[12:32].hex() == address[2:].lower()
    assert int.from_bytes(enc[64:96], "big") == 96
    assert int.from_bytes(enc[96:128], "big") == 34
    c = generate_commitment(address, taskid, cid)
    assert c == keccak256_hex(enc) and len(c) == 66
    a, t, cc = decode(["address", "bytes32", "bytes"], enc)
    assert a == address.lower() and t.hex() == taskid[2:] and cc.hex() == cid[2:]


def test_call_roundtrip_submit_task():
    sig = "submitTask(uint8,address,bytes32,uint256,bytes)"
    payload = json.dumps({"prompt": "arbius test cat"}, separators=(",", ":")).encode()
    data = encode_call(sig, [0, "0x" + "11" * 20, "0x" + "22" * 32, 5, payload])
    v, owner, model, fee, inp = decode_call(sig, data)
    assert (v, fee, inp) == (0, 5, payload) and model == b"\x22" * 32


@given(st.integers(min_value=0, max_value=2**256 - 1))
@settings(max_examples=100, deadline=None)
def test_seed_property(v):
    s = taskid2seed(hex(v))
    assert s == v % 0x1FFFFFFFFFFFF0 and s < 2**53
    assert taskid2seed(v.to_bytes(32, "big")) == s


def test_hash_task_is_deterministic_and_sender_bound():
    cid = IPFS_GOLDEN["d"]
    a = hash_task("0x" + "01" * 32, 0, cid, "0x" + "aa" * 20, "0x" + "00" * 32)
    b = hash_task("0x" + "01" * 32, 0, cid, "0x" + "ab" * 20, "0x" + "00" * 32)
    assert a != b and len(a) == 66


# ---------------- hydration ----------------
def test_hydrate_defaults_and_required():
    t = load_template("anythingv3")
    r = hydrate_input({"prompt": "cat", "negative_prompt": "dog"}, t)
    assert not r.err
    assert r.input["width"] == 768 and r.input["num_inference_steps"] == 20
    assert r.input["guidance_scale"] == 12 and r.input["scheduler"] == "DPMSolverMultistep"
    assert list(r.input) == [row["variable"] for row in t["input"]]
    r = hydrate_input({"prompt": "cat"}, t)
    assert r.err and "negative_prompt" in r.errmsg


def test_hydrate_reference_quirks():
    t = load_template("anythingv3")
    base = {"prompt": "cat", "negative_prompt": ""}
    # B.1 decimal rejects non-integers
    assert hydrate_input({**base, "guidance_scale": 7.5}, t).err
    assert not hydrate_input({**base, "guidance_scale": 7}, t).err
    assert not hydrate_input({**base, "guidance_scale": 7.0}, t).err
    # B.2 max never enforced, min is
    assert not hydrate_input({**base, "num_inference_steps": 100000}, t).err
    assert hydrate_input({**base, "num_inference_steps": 0}, t).err
    # B.3 int limited to int32
    assert hydrate_input({**base, "num_inference_steps": 2**31}, t).err
    # enum
    assert hydrate_input({**base, "width": 777}, t).err
    assert not hydrate_input({**base, "width": 512}, t).err
    assert hydrate_input({**base, "scheduler": "nope"}, t).err
    assert hydrate_input({**base, "prompt": 3}, t).err
    assert hydrate_input({**base, "width": "512"}, t).err
    # zeroscope decimal default 17.5 passes because defaults are not validated
    z = hydrate_input({"prompt": "x", "negative_prompt": "y"}, load_template("zeroscopev2xl"))
    assert not z.err and z.input["guidance_scale"] == 17.5
    # file type unchecked
    r = hydrate_input({"input_video": 123}, load_template("robust_video_matting"))
    assert not r.err and r.input["output_type"] == ""


def test_model_filter():
    t = load_template("kandinsky2")
    models = [ModelEntry(id="0xabc", template=t, filters=[MiningFilter(minfee=10, mintime=5)])]
    ok = check_model_filter(models, model="0xabc", now=100, fee=10, blocktime=90, owner="0x1")
    assert ok.model_enabled and ok.filter_passed and ok.model_template is t
    assert not check_model_filter(models, model="0xabc", now=100, fee=9, blocktime=90, owner="0x1").filter_passed
    assert not check_model_filter(models, model="0xabc", now=100, fee=10, blocktime=97, owner="0x1").filter_passed
    r = check_model_filter(models, model="0xdef", now=100, fee=10, blocktime=0, owner="0x1")
    assert not r.model_enabled and r.model_template is None
    models = [ModelEntry(id="0xabc", template=t, filters=[MiningFilter(owner="0xAA")])]
    assert check_model_filter(models, model="0xabc", now=1, fee=0, blocktime=0, owner="0xaa").filter_passed
    assert not check_model_filter(models, model="0xabc", now=1, fee=0, blocktime=0, owner="0xbb").filter_passed


def test_example_config_loads_unchanged():
    c = load_mining_config(DATA / "MiningConfig.example.json")
    assert c.db_path == "db.sqlite" and c.rpc.port == 8335 and c.automine.delay == 60
    assert c.ml.strategy == "cog" and c.ipfs.strategy == "pinata"
    mid = STATIC_CHAIN_CONFIG["models"]["kandinsky2"]["id"]
    assert c.ml.cog[mid]["url"] == "CHANGEME" and "//" not in c.ml.cog
    assert c.scheduler.poll_ms == 100
