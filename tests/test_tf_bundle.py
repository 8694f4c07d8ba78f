"""CPU tests of the TensorFlow checkpoint-V2 reader / writer (cloud_server_b200/tf_bundle.py).

PARITY UNPINNED against TensorFlow itself (not installable here, the reference ships no checkpoint): these tests pin the
published pieces of the format -- CRC-32C known answers (RFC 3720 B.4), the masked-crc rule, the table magic and
footer, protobuf field numbers -- and the round trip, including hand-built tables whose block layout differs from the
one the writer produces (small blocks, prefix-compressed keys), as a reader must accept."""
import struct

import numpy as np
import pytest


@pytest.fixture(scope="module")
def tb(built_lib):
    from cloud_server_b200 import tf_bundle
    return tf_bundle


def test_crc32c_known_answers(tb):
    # RFC 3720 appendix B.4 test vectors
    assert tb.crc32c(bytes(32)) == 0x8A9136AA
    assert tb.crc32c(b"\xff" * 32) == 0x62A8AB43
    assert tb.crc32c(bytes(range(32))) == 0x46DD794E
    assert tb.crc32c(bytes(range(31, -1, -1))) == 0x113FDB5C
    assert tb.crc32c(b"123456789") == 0xE3069283
    # streaming: continuing from a previous value equals one pass; unaligned starts
    data = np.random.default_rng(0).integers(0, 256, 10007, np.uint8).tobytes()
    assert tb.crc32c(data[5000:], tb.crc32c(data[:5000])) == tb.crc32c(data)
    assert tb.crc32c(data[3:]) == tb.crc32c(bytes(data[3:]))
    # LevelDB / TensorFlow mask: rotate right 15, add 0xa282ead8
    c = tb.crc32c(b"foo")
    assert tb.masked_crc(b"foo") == ((((c >> 15) | (c << 17)) + 0xa282ead8) & 0xFFFFFFFF)


def test_table_round_trip_and_layout(tb, tmp_path):
    items = [(b"", b"hdr")] + [(("Variable_%d" % i).encode(), bytes([i]) * (i + 1)) for i in sorted(range(40), key=lambda i: "Variable_%d" % i)]
    p = str(tmp_path / "t.index")
    tb.write_table(p, items, block_size=64)           # tiny blocks: many data blocks, a real index block
    raw = open(p, "rb").read()
    assert struct.unpack("<Q", raw[-8:])[0] == 0xdb4775248b80fb57 and len(raw) > 48
    assert tb.read_table(p) == items
    tb.write_table(p, items)                          # default: one 256 KB block
    assert tb.read_table(p) == items
    with pytest.raises(ValueError):
        tb.write_table(p, [(b"b", b""), (b"a", b"")])
    # a damaged block is detected by its crc
    bad = bytearray(open(p, "rb").read())
    bad[10] ^= 0x40
    open(p, "wb").write(bytes(bad))
    with pytest.raises(ValueError):
        tb.read_table(p)


def test_bundle_round_trip_with_reference_names(tb, tmp_path):
    rng = np.random.default_rng(1)
    tensors = {"Variable": rng.standard_normal((5, 5, 1, 32)).astype(np.float32), "Variable_1": np.full(32, 0.1, np.float32),
               "Variable_2": rng.standard_normal((3136, 16)).astype(np.float32), "Variable_3": np.array(1234, np.int32),
               "Variable/Adagrad": np.full((5, 5, 1, 32), 0.1, np.float32), "Variable_10": np.arange(6, dtype=np.int64).reshape(2, 3)}
    prefix = str(tmp_path / "train_model" / "model.ckpt-1234")
    tb.write_bundle(prefix, tensors)
    import os
    assert sorted(os.listdir(tmp_path / "train_model")) == ["model.ckpt-1234.data-00000-of-00001", "model.ckpt-1234.index", "model.ckpt-1234.meta"]
    keys = [k for k, _ in tb.read_table(prefix + ".index")]
    assert keys == [b"", b"Variable", b"Variable/Adagrad", b"Variable_1", b"Variable_10", b"Variable_2", b"Variable_3"]   # bytewise order
    back = tb.read_bundle(prefix)
    assert set(back) == set(tensors)
    for k in tensors:
        assert back[k].dtype == tensors[k].dtype and back[k].shape == tensors[k].shape and np.array_equal(back[k], tensors[k])
    # data file = the tensors back to back in key order
    size = sum(tensors[k.decode()].nbytes for k in keys[1:])
    assert os.path.getsize(prefix + ".data-00000-of-00001") == size
    # entry protobuf: dtype = 1 (field 1), shape (field 2), offset (4), size (5), crc32c (6, fixed32)
    entry = dict(tb.read_table(prefix + ".index"))[b"Variable_1"]
    e = tb._parse_entry(entry)
    assert e["dtype"] == 1 and e["shape"] == [32] and e["size"] == 128 and e["offset"] == 2 * 800 * 4
    assert e["crc32c"] == tb.masked_crc(tensors["Variable_1"].tobytes())
    # flipped data byte -> crc failure on read
    with open(prefix + ".data-00000-of-00001", "r+b") as f:
        f.seek(17)
        b = f.read(1)
        f.seek(17)
        f.write(bytes([b[0] ^ 1]))
    with pytest.raises(ValueError):
        tb.read_bundle(prefix)
    assert tb.read_bundle(prefix, check_crc=False)["Variable_1"].shape == (32,)


def test_reader_accepts_prefix_compressed_hand_built_index(tb, tmp_path):
    """An index written by someone else: 2-entry restart interval, shared key prefixes, header with explicit fields."""
    prefix = str(tmp_path / "m.ckpt-7")
    a = np.arange(12, dtype=np.float32).reshape(3, 4)
    g = np.array(7, np.int32)
    open(prefix + ".data-00000-of-00001", "wb").write(a.tobytes() + g.tobytes())
    hdr = b"\x08\x01\x10\x00\x1a\x02\x08\x01"                       # num_shards 1, endianness LITTLE (explicit), version{producer 1}
    ea = tb._entry_proto(1, (3, 4), 0, 48, tb.masked_crc(a.tobytes()))
    eg = b"\x08\x03\x12\x00\x18\x00\x20\x30\x28\x04\x35" + struct.pack("<I", tb.masked_crc(g.tobytes()))   # explicit shard_id 0, offset 48
    blk = tb._BlockBuilder(2)
    for k, v in ((b"", hdr), (b"Variable", ea), (b"Variable_1", eg)):
        blk.add(k, v)
    assert blk.buf.count(b"Variable") == 2            # second restart group starts with the full key; in-group keys are shortened
    with open(prefix + ".index", "wb") as f:
        index = tb._BlockBuilder(1)
        index.add(b"Variable_2", tb._write_block(f, blk.finish()))      # separator key > last key, as FindShortSuccessor may give
        mh = tb._write_block(f, tb._BlockBuilder(1).finish())
        ih = tb._write_block(f, index.finish())
        f.write((mh + ih).ljust(40, b"\x00") + struct.pack("<Q", tb.TABLE_MAGIC))
    back = tb.read_bundle(prefix)
    assert np.array_equal(back["Variable"], a) and int(back["Variable_1"]) == 7
