"""CPU tests of the safetensors reader / writer (vllm-rust_b200/checkpoint.py): the wire format the reference mmaps into
its VarBuilder (loader.rs:189-198, 418-444).  Fixed bytes first (a file assembled by hand, so the reader is not only
tested against its own writer), then round trips, the shard index, and the error paths."""
import json
import struct

import numpy as np
import pytest
import torch


def hand_made(path):
    """header + payload written byte by byte, per the published layout."""
    a = np.arange(6, dtype="<i4").reshape(2, 3)
    b = np.array([0x3F80, 0xC000], dtype="<u2")                      # bf16 bits of 1.0, -2.0
    c = np.array([1.5, -0.25], dtype="<f2")
    header = {"__metadata__": {"format": "pt"},
              "m.qweight": {"dtype": "I32", "shape": [2, 3], "data_offsets": [0, 24]},
              "m.bias": {"dtype": "BF16", "shape": [2], "data_offsets": [24, 28]},
              "m.scales": {"dtype": "F16", "shape": [1, 2], "data_offsets": [28, 32]},
              "m.u": {"dtype": "U32", "shape": [1], "data_offsets": [32, 36]}}
    hj = json.dumps(header).encode()
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(hj)) + hj + a.tobytes() + b.tobytes() + c.tobytes() + struct.pack("<I", 0xFFFFFFFE))


def test_reads_a_hand_made_file(product, tmp_path):
    p = tmp_path / "model.safetensors"
    hand_made(p)
    ck = product.SafetensorsCheckpoint(tmp_path)
    assert sorted(ck) == ["m.bias", "m.qweight", "m.scales", "m.u"] and len(ck) == 4
    assert torch.equal(ck["m.qweight"], torch.arange(6, dtype=torch.int32).reshape(2, 3))
    assert ck["m.bias"].dtype is torch.bfloat16 and ck["m.bias"].tolist() == [1.0, -2.0]
    assert ck["m.scales"].dtype is torch.float16 and ck["m.scales"].tolist() == [[1.5, -0.25]]
    assert ck["m.u"].dtype is torch.int32 and ck["m.u"].item() == -2          # U32 handed out as the same bits
    assert ck.shape("m.qweight") == (2, 3) and ck.dtype_tag("m.u") == "U32"
    assert ck.quantized_prefixes() == ["m"]
    assert ck.get("nope") is None and "nope" not in ck and "m.bias" in ck
    with pytest.raises(KeyError):
        ck["nope"]
    assert product.SafetensorsCheckpoint(p).fingerprint() == ck.fingerprint()


def test_round_trip_and_shards(product, tmp_path):
    g = torch.Generator().manual_seed(0)
    t = {"a.qweight": torch.randint(-2**31, 2**31 - 1, (64, 16), dtype=torch.int32, generator=g),
         "a.scales": torch.rand(2, 128, generator=g).to(torch.float16),
         "a.bias": torch.randn(128, generator=g).to(torch.bfloat16),
         "a.g_idx": torch.randint(0, 2, (64,), dtype=torch.int32, generator=g),
         "n.weight": torch.randn(4, 4, generator=g), "empty": torch.zeros(0, 3)}
    product.save_safetensors(t, tmp_path / "model.safetensors")
    one = product.SafetensorsCheckpoint(tmp_path)
    for k, v in t.items():
        assert one[k].dtype == v.dtype and torch.equal(one[k], v), k
    d = tmp_path / "sharded"
    product.save_sharded(t, d, 3)
    many = product.SafetensorsCheckpoint(d)
    assert sorted(many) == sorted(t)
    for k, v in t.items():
        assert torch.equal(many[k], v)
    assert many.fingerprint() != one.fingerprint()
    assert many.dense_prefixes() == ["n"]
    # the index must agree with the shards
    idx = json.loads((d / "model.safetensors.index.json").read_text())
    idx["weight_map"]["a.bias"] = "model-00001-of-00003.safetensors" if idx["weight_map"]["a.bias"] != "model-00001-of-00003.safetensors" else "model-00002-of-00003.safetensors"
    (d / "model.safetensors.index.json").write_text(json.dumps(idx))
    with pytest.raises(product.W4A16Error, match="index maps"):
        product.SafetensorsCheckpoint(d)
    (d / "model.safetensors.index.json").write_text('{"metadata": {}}')
    with pytest.raises(product.W4A16Error, match="missing weight_map"):
        product.SafetensorsCheckpoint(d)


def test_natural_shard_order(product, tmp_path):
    for i in (10, 2, 1):
        product.save_safetensors({f"t{i}": torch.tensor([i])}, tmp_path / f"part-{i}.safetensors")
    ck = product.SafetensorsCheckpoint(tmp_path)
    assert [f.path.rsplit("-", 1)[1] for f in ck._files] == ["1.safetensors", "2.safetensors", "10.safetensors"]


def test_malformed_files_are_refused(product, tmp_path):
    def write(name, header, payload=b"", raw_header=None):
        hj = raw_header if raw_header is not None else json.dumps(header).encode()
        (tmp_path / name).write_bytes(struct.pack("<Q", len(hj)) + hj + payload)
        return tmp_path / name
    E = product.W4A16Error
    (tmp_path / "short.safetensors").write_bytes(b"\1\2\3")
    with pytest.raises(E, match="shorter than 8"):
        product.SafetensorsCheckpoint(tmp_path / "short.safetensors")
    (tmp_path / "len.safetensors").write_bytes(struct.pack("<Q", 1 << 40) + b"{}")
    with pytest.raises(E, match="header length"):
        product.SafetensorsCheckpoint(tmp_path / "len.safetensors")
    with pytest.raises(E, match="not JSON"):
        product.SafetensorsCheckpoint(write("json.safetensors", None, raw_header=b"{nope"))
    with pytest.raises(E, match="unsupported dtype"):
        product.SafetensorsCheckpoint(write("dt.safetensors", {"x": {"dtype": "F8_E4M3", "shape": [1], "data_offsets": [0, 1]}}, b"\0"))
    with pytest.raises(E, match="do not match shape"):
        product.SafetensorsCheckpoint(write("off.safetensors", {"x": {"dtype": "I32", "shape": [2], "data_offsets": [0, 4]}}, b"\0" * 8))
    with pytest.raises(E, match="do not match shape"):     # runs past the end of the file
        product.SafetensorsCheckpoint(write("end.safetensors", {"x": {"dtype": "I32", "shape": [4], "data_offsets": [0, 16]}}, b"\0" * 8))
    with pytest.raises(E, match="malformed header entry"):
        product.SafetensorsCheckpoint(write("ent.safetensors", {"x": {"dtype": "I32"}}))
    with pytest.raises(E, match="no safetensors files"):
        product.SafetensorsCheckpoint(tmp_path / "nothing_here")
    a = write("dup1.safetensors", {"x": {"dtype": "U8", "shape": [1], "data_offsets": [0, 1]}}, b"\0")
    b = write("dup2.safetensors", {"x": {"dtype": "U8", "shape": [1], "data_offsets": [0, 1]}}, b"\0")
    with pytest.raises(E, match="appears in both"):
        product.SafetensorsCheckpoint([a, b])


def test_loader_over_a_checkpoint_validates_shapes_on_the_cpu(product, tmp_path):
    """The loaders' shape errors (weight_loader.rs:596-714 wording) fire before anything touches the device."""
    t = {"l.qweight": torch.zeros(256, 8, dtype=torch.int32), "l.qzeros": torch.zeros(2, 8, dtype=torch.int32),
         "l.scales": torch.zeros(2, 64, dtype=torch.float16)}
    product.save_safetensors(t, tmp_path / "model.safetensors")
    ck = product.SafetensorsCheckpoint(tmp_path)
    det = product.DetectedQuantConfig(product.QuantizationMethod.AWQ, 4, 128, None, {"modules_to_not_convert": ["skipme"]})
    loader = product.create_weight_loader_with_params(ck, det, "cuda:0", cache_dir=tmp_path / "c")
    with pytest.raises(product.W4A16Error, match=r"AWQ qweight load failed for l with expected shape \[256, 16\]"):
        loader.load_linear("l", 256, 128, False)
    with pytest.raises(product.W4A16Error, match="missing tensor"):
        loader.load_linear("other", 256, 64, False)
    with pytest.raises(product.W4A16Error, match="modules_to_not_convert"):
        loader.load_linear("x.skipme", 256, 64, False)
    with pytest.raises(product.W4A16Error, match="one out_features entry per prefix"):
        loader.load_fused(["l"], 256, [32, 32], False)
