"""TensorFlow checkpoint "V2" bundles (what tf.train.Saver writes and restores) -- reader and writer, no TensorFlow.

The reference saves its models with ``tf.train.Saver()`` driven by the Supervisor
(apps/construction/util/construct_distribute.py:375, :385-392) and serves them with ``saver.restore(sess, save_path)``
(construct_inference.py:336-339).  This module reads and writes that on-disk format so that a model trained by the old
platform can be served by this library and vice versa:

  <prefix>.index                an SSTable (the LevelDB table format TensorFlow embeds in tensorflow/core/lib/io):
                                key "" -> BundleHeaderProto, key <variable name> -> BundleEntryProto (dtype, shape,
                                shard, offset, size, masked crc32c of the bytes)
  <prefix>.data-00000-of-00001  the tensors' raw little-endian bytes, back to back, in key order
  <prefix>.meta                 the MetaGraphDef.  Neither restore path above reads it (both rebuild the graph in Python),
                                but get_model() (construct_inference.py:188-203) lists *.meta to find the newest step:
                                an empty marker file is written.

Variable names: the reference creates every variable as an unnamed ``tf.Variable`` (:57-87), so TensorFlow names them
``Variable, Variable_1, ...`` in creation order = the order of this library's flat parameter buffer; ``global_step =
tf.Variable(0)`` (:370, int32) comes next, and AdagradOptimizer adds one ``<name>/Adagrad`` slot per trained variable.

PARITY UNPINNED: there is no TensorFlow in this image and the reference ships no checkpoint, so there is no fixture to
pin this against; the writer follows the published formats (table_builder.cc / tensor_bundle.proto) and the reader
accepts every block layout the format allows (prefix-compressed keys, any block size, crc checked).  Snappy-compressed
blocks (never produced by BundleWriter) are refused.
"""
import ctypes as C
import os
import struct

import numpy as np

from . import _lib

TABLE_MAGIC = 0xdb4775248b80fb57
MASK_DELTA = 0xa282ead8
DT_FLOAT, DT_INT32, DT_INT64 = 1, 3, 9
_NP_OF = {DT_FLOAT: np.dtype("<f4"), DT_INT32: np.dtype("<i4"), DT_INT64: np.dtype("<i8")}
_DT_OF = {np.dtype("float32"): DT_FLOAT, np.dtype("int32"): DT_INT32, np.dtype("int64"): DT_INT64}


def crc32c(data, crc=0):
    """CRC-32C (Castagnoli) of a bytes-like / contiguous numpy array, computed by the C library (host code, no GPU)."""
    buf = np.frombuffer(data, np.uint8) if not isinstance(data, np.ndarray) else np.ascontiguousarray(data).view(np.uint8).reshape(-1)
    if buf.size == 0:
        return crc
    return int(_lib.lib().nc_crc32c(buf.ctypes.data_as(C.c_void_p), C.c_size_t(buf.size), C.c_uint32(crc)))


def masked_crc(data):
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + MASK_DELTA) & 0xFFFFFFFF


# ---------------------------------------------------------------------------------------------- varints / protobuf
def _varint(v):
    out = bytearray()
    v &= (1 << 64) - 1
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _read_varint(b, i):
    v, shift = 0, 0
    while True:
        c = b[i]
        i += 1
        v |= (c & 0x7F) << shift
        if c < 0x80:
            return v, i
        shift += 7


def _pb_fields(b):
    """yield (field_number, wire_type, value) of one serialized message"""
    i = 0
    while i < len(b):
        tag, i = _read_varint(b, i)
        f, wt = tag >> 3, tag & 7
        if wt == 0:
            v, i = _read_varint(b, i)
        elif wt == 1:
            v = b[i:i + 8]
            i += 8
        elif wt == 2:
            n, i = _read_varint(b, i)
            v = b[i:i + n]
            i += n
        elif wt == 5:
            v = b[i:i + 4]
            i += 4
        else:
            raise ValueError("unsupported protobuf wire type %d" % wt)
        yield f, wt, v


def _header_proto(num_shards=1):
    # BundleHeaderProto{num_shards = 1; endianness = 2 (LITTLE = 0: default, omitted); version = 3 {producer = 1}}
    version = b"\x08\x01"
    return b"\x08" + _varint(num_shards) + b"\x1a" + _varint(len(version)) + version


def _entry_proto(dtype, shape, offset, size, crc):
    dims = b"".join(b"\x12" + _varint(len(d)) + d for d in (b"\x08" + _varint(int(s)) for s in shape))   # TensorShapeProto.dim = 2 {size = 1}
    out = b"\x08" + _varint(dtype) + b"\x12" + _varint(len(dims)) + dims
    if offset:
        out += b"\x20" + _varint(offset)                      # shard_id = 3 stays 0 (omitted); offset = 4
    out += b"\x28" + _varint(size) + b"\x35" + struct.pack("<I", crc)   # size = 5, crc32c = 6 (fixed32)
    return out


def _parse_entry(b):
    e = {"dtype": 0, "shape": [], "shard_id": 0, "offset": 0, "size": 0, "crc32c": None, "slices": 0}
    for f, wt, v in _pb_fields(b):
        if f == 1:
            e["dtype"] = v
        elif f == 2:
            for f2, _, v2 in _pb_fields(v):
                if f2 == 2:
                    size = 0
                    for f3, _, v3 in _pb_fields(v2):
                        if f3 == 1:
                            size = v3 if v3 < (1 << 63) else v3 - (1 << 64)
                    e["shape"].append(size)
        elif f == 3:
            e["shard_id"] = v
        elif f == 4:
            e["offset"] = v
        elif f == 5:
            e["size"] = v
        elif f == 6:
            e["crc32c"] = struct.unpack("<I", v)[0]
        elif f == 7:
            e["slices"] += 1
    return e


# ---------------------------------------------------------------------------------------------- SSTable
class _BlockBuilder:
    def __init__(self, restart_interval):
        self.ri, self.buf, self.restarts, self.count, self.last = restart_interval, bytearray(), [0], 0, b""

    def add(self, key, value):
        shared = 0
        if self.count % self.ri == 0 and self.count:
            self.restarts.append(len(self.buf))
        elif self.count:
            m = min(len(key), len(self.last))
            while shared < m and key[shared] == self.last[shared]:
                shared += 1
        self.buf += _varint(shared) + _varint(len(key) - shared) + _varint(len(value)) + key[shared:] + value
        self.last, self.count = key, self.count + 1

    def finish(self):
        return bytes(self.buf) + b"".join(struct.pack("<I", r) for r in self.restarts) + struct.pack("<I", len(self.restarts))


def _write_block(f, contents):
    """block + trailer (type 0 = uncompressed, masked crc32c over contents + type); returns its BlockHandle bytes"""
    off = f.tell()
    f.write(contents)
    f.write(b"\x00" + struct.pack("<I", masked_crc(contents + b"\x00")))
    return _varint(off) + _varint(len(contents))


def write_table(path, items, block_size=262144):
    """items: list of (key bytes, value bytes), keys strictly increasing"""
    if any(b[0] <= a[0] for a, b in zip(items, items[1:])):
        raise ValueError("table keys must be strictly increasing")
    with open(path, "wb") as f:
        index = _BlockBuilder(1)
        blk, last_key = _BlockBuilder(16), None
        for k, v in items:
            blk.add(k, v)
            last_key = k
            if len(blk.buf) >= block_size:
                index.add(last_key, _write_block(f, blk.finish()))     # the block's own last key is a valid separator
                blk = _BlockBuilder(16)
        if blk.count or last_key is None:
            index.add(last_key or b"", _write_block(f, blk.finish()))
        meta_handle = _write_block(f, _BlockBuilder(1).finish())        # empty metaindex block
        index_handle = _write_block(f, index.finish())
        footer = meta_handle + index_handle
        f.write(footer + b"\x00" * (40 - len(footer)) + struct.pack("<Q", TABLE_MAGIC))


def _read_block(buf, handle_bytes, i=0):
    off, i = _read_varint(handle_bytes, i)
    size, i = _read_varint(handle_bytes, i)
    contents, typ = buf[off:off + size], buf[off + size]
    crc = struct.unpack("<I", buf[off + size + 1:off + size + 5])[0]
    if masked_crc(bytes(contents) + bytes([typ])) != crc:
        raise ValueError("table block at offset %d fails its crc32c" % off)
    if typ != 0:
        raise ValueError("compressed table block (type %d): BundleWriter never writes these" % typ)
    return bytes(contents), i


def _block_entries(contents):
    nrestarts = struct.unpack("<I", contents[-4:])[0]
    end = len(contents) - 4 - 4 * nrestarts
    i, key = 0, b""
    while i < end:
        shared, i = _read_varint(contents, i)
        unshared, i = _read_varint(contents, i)
        vlen, i = _read_varint(contents, i)
        key = key[:shared] + contents[i:i + unshared]
        i += unshared
        yield key, contents[i:i + vlen]
        i += vlen


def read_table(path):
    buf = open(path, "rb").read()
    if len(buf) < 48 or struct.unpack("<Q", buf[-8:])[0] != TABLE_MAGIC:
        raise ValueError("%s is not a TensorFlow checkpoint index (bad table magic)" % path)
    footer = buf[-48:-8]
    _, i = _read_block(buf, footer)                 # metaindex (unused)
    index, _ = _read_block(buf, footer, i)
    out = []
    for _, handle in _block_entries(index):
        blk, _ = _read_block(buf, handle)
        out.extend(_block_entries(blk))
    return out


# ---------------------------------------------------------------------------------------------- bundles
def write_bundle(prefix, tensors):
    """tensors: {name: numpy array (float32 / int32 / int64)} -> <prefix>.index + <prefix>.data-00000-of-00001 (+ .meta marker)"""
    os.makedirs(os.path.dirname(prefix) or ".", exist_ok=True)
    items = [(b"", _header_proto(1))]
    off = 0
    with open(prefix + ".data-00000-of-00001", "wb") as f:
        for name in sorted(tensors, key=lambda s: s.encode()):
            a = np.asarray(tensors[name])          # (ascontiguousarray would turn a scalar into shape (1,))
            if a.dtype not in _DT_OF:
                raise ValueError("unsupported dtype %s for %s" % (a.dtype, name))
            raw = a.astype(a.dtype.newbyteorder("<"), copy=False).tobytes(order="C")
            f.write(raw)
            items.append((name.encode(), _entry_proto(_DT_OF[a.dtype], a.shape, off, len(raw), masked_crc(raw))))
            off += len(raw)
    write_table(prefix + ".index", items)
    open(prefix + ".meta", "ab").close()
    return prefix


def read_bundle(prefix, check_crc=True):
    """-> {name: numpy array}; raises ValueError on a damaged or unsupported bundle"""
    entries = read_table(prefix + ".index")
    if not entries or entries[0][0] != b"":
        raise ValueError("checkpoint index has no header entry")
    num_shards = 1
    for f, _, v in _pb_fields(entries[0][1]):
        if f == 1:
            num_shards = v
        if f == 2 and v != 0:
            raise ValueError("big-endian checkpoint")
    data = {}
    out = {}
    for key, val in entries[1:]:
        e = _parse_entry(val)
        if e["slices"]:
            raise ValueError("partitioned variable %r: not produced by the reference's Saver" % key.decode())
        if e["dtype"] not in _NP_OF:
            raise ValueError("variable %r has unsupported dtype %d" % (key.decode(), e["dtype"]))
        sid = e["shard_id"]
        if sid not in data:
            data[sid] = np.memmap("%s.data-%05d-of-%05d" % (prefix, sid, num_shards), dtype=np.uint8, mode="r")
        raw = np.asarray(data[sid][e["offset"]:e["offset"] + e["size"]])
        if raw.size != e["size"]:
            raise ValueError("variable %r runs past the end of its data shard" % key.decode())
        if check_crc and e["crc32c"] is not None and masked_crc(raw) != e["crc32c"]:
            raise ValueError("variable %r fails its crc32c" % key.decode())
        out[key.decode()] = raw.view(_NP_OF[e["dtype"]]).reshape(e["shape"]).copy()
    return out


# ---------------------------------------------------------------------------------------------- the reference's naming
def variable_name(i):
    return "Variable" if i == 0 else "Variable_%d" % i


def save_net(prefix, net, opt_id=0):
    """Write `net` (cloud_server_b200.net.Net) as the checkpoint the reference's Saver would write at this step: variables
    in creation order, global_step as the next unnamed int32 variable, Adagrad accumulators as <name>/Adagrad slots."""
    flat = net.get_params()
    tensors = {}
    for i, (off, size, shape) in enumerate(net.tensors):
        tensors[variable_name(i)] = flat[off:off + size].reshape(shape)
    tensors[variable_name(len(net.tensors))] = np.array(net.global_step, np.int32)
    if opt_id == 0:
        acc = net.get_state(0)
        for i, (off, size, shape) in enumerate(net.tensors):
            tensors[variable_name(i) + "/Adagrad"] = acc[off:off + size].reshape(shape)
    elif opt_id == 2:
        # tf.train.AdamOptimizer (TF 1.x): slots <name>/Adam (m) and <name>/Adam_1 (v); beta1_power / beta2_power are
        # non-slot variables that start at beta and are multiplied by beta after every step
        mv = net.get_state(2)
        for i, (off, size, shape) in enumerate(net.tensors):
            tensors[variable_name(i) + "/Adam"] = mv[off:off + size].reshape(shape)
            tensors[variable_name(i) + "/Adam_1"] = mv[net.param_count + off:net.param_count + off + size].reshape(shape)
        tensors["beta1_power"] = np.array(0.9 ** (net.global_step + 1), np.float32)
        tensors["beta2_power"] = np.array(0.999 ** (net.global_step + 1), np.float32)
    return write_bundle(prefix, tensors)


def load_net(prefix, net):
    """Restore what saver.restore(sess, prefix) restores: every variable of the rebuilt graph by name; shapes must match.
    Adagrad slots and the step counter are restored when present.  Returns the global step (0 if absent)."""
    t = read_bundle(prefix)
    flat = np.empty(net.param_count, np.float32)
    for i, (off, size, shape) in enumerate(net.tensors):
        name = variable_name(i)
        if name not in t:
            raise ValueError("Key %s not found in checkpoint" % name)            # TF: NotFoundError
        if tuple(t[name].shape) != tuple(shape):
            raise ValueError("checkpoint tensor %s has shape %s, the graph wants %s" % (name, t[name].shape, shape))
        flat[off:off + size] = t[name].reshape(-1)
    net.set_params(flat)
    gs = t.get(variable_name(len(net.tensors)))
    step = int(gs) if gs is not None and gs.ndim == 0 else 0
    if all(variable_name(i) + "/Adagrad" in t for i in range(len(net.tensors))):
        acc = np.empty(net.param_count, np.float32)
        for i, (off, size, shape) in enumerate(net.tensors):
            acc[off:off + size] = t[variable_name(i) + "/Adagrad"].reshape(-1)
        net.set_state(0, acc, step)
    elif all(variable_name(i) + "/Adam" in t and variable_name(i) + "/Adam_1" in t for i in range(len(net.tensors))):
        mv = np.empty(2 * net.param_count, np.float32)
        for i, (off, size, shape) in enumerate(net.tensors):
            mv[off:off + size] = t[variable_name(i) + "/Adam"].reshape(-1)
            mv[net.param_count + off:net.param_count + off + size] = t[variable_name(i) + "/Adam_1"].reshape(-1)
        net.set_state(2, mv, step)
    else:
        net.set_state(0, None, step)
    return step
