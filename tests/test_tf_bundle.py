"""TF checkpoint V2 (tensor bundle) writer / reader: round trip, table structure, checksums."""
import os
import struct

import pytest
import torch

from dtg_b200.session import tf_bundle as tb


def test_crc32c_known_vectors():
    assert tb.crc32c(b"123456789") == 0xE3069283          # RFC 3720 check value
    assert tb.crc32c(b"\x00" * 32) == 0x8A9136AA
    big = bytes(range(256)) * 64
    py = tb._TABLE and None
    saved, tb._native = tb._native, False                 # pure-Python path
    try:
        py = tb.crc32c(big)
    finally:
        tb._native = saved
    assert tb.crc32c(big) == py                           # native slicing-by-8 agrees (if built)
    m = tb.masked_crc32c(b"abc")
    assert tb._unmask(m) == tb.crc32c(b"abc")


def test_bundle_round_trip_and_layout(tmp_path):
    torch.manual_seed(0)
    prefix = str(tmp_path / "model.ckpt-60")
    tensors = {"g/Variable": torch.randn(2), "g/Variable/Adagrad": torch.full((2,), 0.1),
               "g/Variable_1": torch.randn(2), "global_step": torch.tensor(60),
               "wte": torch.randn(300, 64).bfloat16(), "flag": torch.tensor([True, False]),
               "big": torch.randn(400, 400)}
    for i in range(200):                                  # several 4 KB index blocks
        tensors["layer_%03d/weights" % i] = torch.full((3,), float(i))
    tb.write_bundle(prefix, tensors)
    assert sorted(os.listdir(tmp_path)) == ["model.ckpt-60.data-00000-of-00001", "model.ckpt-60.index"]
    assert tb.is_bundle(prefix)
    idx = open(prefix + ".index", "rb").read()
    assert struct.unpack("<Q", idx[-8:])[0] == 0xdb4775248b80fb57        # leveldb table magic
    items = tb._read_table(prefix + ".index")
    keys = [k for k, _ in items]
    assert keys[0] == b"" and keys == sorted(keys) and len(keys) == len(tensors) + 1
    header = tb._parse_proto(items[0][1])
    assert header[1] == [1] and tb._parse_proto(header[3][0])[1] == [1]  # num_shards, version.producer
    listing = tb.list_bundle(prefix)
    assert listing["global_step"] == (torch.int64, ()) and listing["wte"] == (torch.bfloat16, (300, 64))
    out = tb.read_bundle(prefix)
    for k, v in tensors.items():
        assert out[k].dtype == v.dtype and torch.equal(out[k], v), k
    assert set(tb.read_bundle(prefix, names=["big"])) == {"big"}
    # data file is the plain concatenation in key order
    assert os.path.getsize(prefix + ".data-00000-of-00001") == sum(
        v.numel() * v.element_size() for v in tensors.values())


def test_bundle_detects_corruption(tmp_path):
    prefix = str(tmp_path / "m")
    tb.write_bundle(prefix, {"a": torch.arange(1000, dtype=torch.float32)})
    path = prefix + ".data-00000-of-00001"
    raw = bytearray(open(path, "rb").read())
    raw[100] ^= 0xFF
    open(path, "wb").write(bytes(raw))
    with pytest.raises(ValueError, match="checksum"):
        tb.read_bundle(prefix)
    assert tb.read_bundle(prefix, verify=False)["a"].shape == (1000,)


def test_saver_writes_and_restores_tf_layout(dtg, tmp_path):
    a = dtg.Variable(dtg.zeros([2]))
    gs = dtg.Variable(0, dtype=torch.int64, trainable=False, name="global_step")
    sess = dtg.train.Supervisor().prepare_or_wait_for_session()
    a.assign(torch.tensor([1.5, -2.0])); gs.assign(7)
    saver = dtg.train.Saver(write_format="tf_bundle")
    prefix = saver.save(sess, str(tmp_path / "model.ckpt"), global_step=gs)
    assert os.path.exists(prefix + ".index") and prefix.endswith("model.ckpt-7")
    assert dtg.train.latest_checkpoint(str(tmp_path)) == prefix       # TF-syntax state file
    a.assign(torch.zeros(2))
    assert "Variable" in saver.restore(sess, prefix)
    assert a.read_value().tolist() == [1.5, -2.0]
    data = dtg.train.load_checkpoint(prefix)
    assert int(data["global_step"]) == 7
