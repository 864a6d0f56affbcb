"""TEST INFRASTRUCTURE (oracle) for the tf.Example / TFRecord egress.

Only tests/ may import this.  It restates, with the real protobuf runtime, what the reference gets
from TensorFlow: ``tf.train.Example`` (tensorflow/core/example/example.proto + feature.proto,
field numbers restated below [UPSTREAM-RECALLED]) serialised deterministically, and the TFRecord
framing (tensorflow/core/lib/io/record_writer.cc [UPSTREAM-RECALLED]): u64 length, masked CRC-32C
of the length, data, masked CRC-32C of the data; mask(c) = rotr(c, 15) + 0xa282ead8.
Pins: CRC-32C check value 0xE3069283 for b"123456789" (Castagnoli, published); the Example bytes
come from the protobuf library itself.  The reference's own files (reference call sites:
notebooks/03a_TensorFlow_Transform_Advanced.ipynb:445-455, airflow-dags/taxi_utils.py:79-83) are
not available here, so agreement with TensorFlow's writer is **unpinned** beyond that.
"""
import struct

import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

_F = descriptor_pb2.FieldDescriptorProto


def _build_example_class():
    fd = descriptor_pb2.FileDescriptorProto(name="txf_example.proto", package="txf_oracle", syntax="proto3")

    def msg(name):
        m = fd.message_type.add()
        m.name = name
        return m

    def field(m, name, number, ftype, label=_F.LABEL_OPTIONAL, type_name=None, oneof=None):
        f = m.field.add()
        f.name, f.number, f.type, f.label = name, number, ftype, label
        if type_name:
            f.type_name = ".txf_oracle." + type_name
        if oneof is not None:
            f.oneof_index = oneof
        return f

    field(msg("BytesList"), "value", 1, _F.TYPE_BYTES, _F.LABEL_REPEATED)
    field(msg("FloatList"), "value", 1, _F.TYPE_FLOAT, _F.LABEL_REPEATED)   # proto3: packed
    field(msg("Int64List"), "value", 1, _F.TYPE_INT64, _F.LABEL_REPEATED)   # proto3: packed
    feat = msg("Feature")
    feat.oneof_decl.add().name = "kind"
    field(feat, "bytes_list", 1, _F.TYPE_MESSAGE, type_name="BytesList", oneof=0)
    field(feat, "float_list", 2, _F.TYPE_MESSAGE, type_name="FloatList", oneof=0)
    field(feat, "int64_list", 3, _F.TYPE_MESSAGE, type_name="Int64List", oneof=0)
    feats = msg("Features")
    entry = feats.nested_type.add()
    entry.name = "FeatureEntry"
    entry.options.map_entry = True
    field(entry, "key", 1, _F.TYPE_STRING)
    field(entry, "value", 2, _F.TYPE_MESSAGE, type_name="Feature")
    field(feats, "feature", 1, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, type_name="Features.FeatureEntry")
    field(msg("Example"), "features", 1, _F.TYPE_MESSAGE, type_name="Features")
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("txf_oracle.Example"))


Example = _build_example_class()

_CRC_TABLE = []
for _i in range(256):
    _c = _i
    for _ in range(8):
        _c = (_c >> 1) ^ 0x82F63B78 if _c & 1 else _c >> 1
    _CRC_TABLE.append(_c)


def crc32c(data: bytes) -> int:
    c = 0xFFFFFFFF
    for b in data:
        c = _CRC_TABLE[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc(data: bytes) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def make_example(row: dict):
    """row: feature -> None (missing) | float (np.float32) | int | bytes, typed by `kinds` of the caller."""
    ex = Example()
    for name, (kind, value) in row.items():
        f = ex.features.feature[name]
        lst = f.bytes_list if kind == "bytes" else f.float_list if kind == "float32" else f.int64_list
        if value is None:
            lst.SetInParent()  # an empty list of that kind
        elif kind == "bytes":
            lst.value.append(value)
        elif kind == "float32":
            lst.value.append(float(np.float32(value)))
        else:
            lst.value.append(int(value))
    return ex


def frame(data: bytes) -> bytes:
    head = struct.pack("<Q", len(data))
    return head + struct.pack("<I", masked_crc(head)) + data + struct.pack("<I", masked_crc(data))


def encode_rows(rows) -> bytes:
    return b"".join(frame(make_example(r).SerializeToString(deterministic=True)) for r in rows)


def decode_tfrecord(buf: bytes):
    """-> list of Example messages; verifies both CRCs of every record."""
    out, p = [], 0
    while p < len(buf):
        head = buf[p:p + 8]
        (n,) = struct.unpack("<Q", head)
        assert struct.unpack("<I", buf[p + 8:p + 12])[0] == masked_crc(head), "length CRC mismatch"
        data = buf[p + 12:p + 12 + n]
        assert struct.unpack("<I", buf[p + 12 + n:p + 16 + n])[0] == masked_crc(data), "data CRC mismatch"
        ex = Example()
        ex.ParseFromString(data)
        out.append(ex)
        p += 16 + n
    return out
