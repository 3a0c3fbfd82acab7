"""Minimal Avro: binary encoding + Object Container Files, schema-driven, no dependencies.

The job-history files ``tony-portal`` reads are Avro OCFs (null codec, writer schema embedded)
written by the reference's EventHandler through the Avro library
(tony-core/src/main/java/com/linkedin/tony/events/EventHandler.java:97-155; read back in
util/ParserUtils.java:258-285).  No Avro library is installed on the B200 image, so the
subset the history format needs is implemented here: null / boolean / int / long / float /
double / string / bytes, records, enums, arrays, maps, unions and named-type references.
The decoder walks the *embedded writer schema*, so older files (e.g. the reference's sample
``tony-portal/example/tony-history/**.jhist``, which predates some fields) still parse.
"""
from __future__ import annotations

import io
import json
import os
import struct
from typing import Any, BinaryIO, Dict, List, Optional, Tuple

MAGIC = b"Obj\x01"
SYNC_SIZE = 16


# ---- primitives ----------------------------------------------------------------------------
def encode_long(n: int) -> bytes:
    z = (n << 1) ^ (n >> 63)
    out = bytearray()
    while (z & ~0x7F) != 0:
        out.append((z & 0x7F) | 0x80)
        z >>= 7
    out.append(z)
    return bytes(out)


def decode_long(buf: BinaryIO) -> int:
    shift, acc = 0, 0
    while True:
        b = buf.read(1)
        if not b:
            raise EOFError("truncated varint")
        v = b[0]
        acc |= (v & 0x7F) << shift
        if not (v & 0x80):
            break
        shift += 7
    return (acc >> 1) ^ -(acc & 1)


def _read_exact(buf: BinaryIO, n: int) -> bytes:
    data = buf.read(n)
    if len(data) != n:
        raise EOFError("truncated Avro data")
    return data


# ---- schema handling -----------------------------------------------------------------------
class Names:
    """Named-type registry built while walking a schema."""

    def __init__(self):
        self.types: Dict[str, Any] = {}

    def register(self, schema: dict, namespace: Optional[str]) -> None:
        name = schema["name"]
        ns = schema.get("namespace", namespace)
        self.types[name] = schema
        if ns and "." not in name:
            self.types[f"{ns}.{name}"] = schema

    def resolve(self, ref: str) -> Any:
        if ref in self.types:
            return self.types[ref]
        short = ref.rsplit(".", 1)[-1]
        if short in self.types:
            return self.types[short]
        raise KeyError(f"unknown Avro type {ref!r}")


PRIMITIVES = {"null", "boolean", "int", "long", "float", "double", "bytes", "string"}


def _collect(schema: Any, names: Names, namespace: Optional[str] = None) -> None:
    if isinstance(schema, list):
        for s in schema:
            _collect(s, names, namespace)
    elif isinstance(schema, dict):
        t = schema.get("type")
        if t in ("record", "enum", "fixed"):
            names.register(schema, namespace)
            ns = schema.get("namespace", namespace)
            for f in schema.get("fields", []):
                _collect(f["type"], names, ns)
        elif t == "array":
            _collect(schema["items"], names, namespace)
        elif t == "map":
            _collect(schema["values"], names, namespace)
        elif isinstance(t, (dict, list)):
            _collect(t, names, namespace)


def _branch_matches(schema: Any, value: Any, names: Names) -> bool:
    if isinstance(schema, str) and schema not in PRIMITIVES:
        schema = names.resolve(schema)
    if isinstance(schema, str):
        return {
            "null": value is None, "boolean": isinstance(value, bool),
            "int": isinstance(value, int) and not isinstance(value, bool),
            "long": isinstance(value, int) and not isinstance(value, bool),
            "float": isinstance(value, float), "double": isinstance(value, (float, int)) and not isinstance(value, bool),
            "bytes": isinstance(value, (bytes, bytearray)), "string": isinstance(value, str),
        }[schema]
    t = schema.get("type")
    if t == "record":
        return isinstance(value, dict) and (value.get("__type__") in (None, schema["name"])) and \
            all(f["name"] in value or "default" in f for f in schema["fields"])
    if t == "enum":
        return isinstance(value, str) and value in schema["symbols"]
    if t == "array":
        return isinstance(value, (list, tuple))
    if t == "map":
        return isinstance(value, dict)
    return _branch_matches(t, value, names)


def encode(schema: Any, value: Any, names: Names, out: bytearray) -> None:
    if isinstance(schema, list):  # union
        for i, branch in enumerate(schema):
            if _branch_matches(branch, value, names):
                out += encode_long(i)
                encode(branch, value, names, out)
                return
        raise ValueError(f"value {value!r} matches no branch of union {schema}")
    if isinstance(schema, str):
        if schema not in PRIMITIVES:
            return encode(names.resolve(schema), value, names, out)
        if schema == "null":
            return
        if schema == "boolean":
            out.append(1 if value else 0)
        elif schema in ("int", "long"):
            out += encode_long(int(value))
        elif schema == "float":
            out += struct.pack("<f", float(value))
        elif schema == "double":
            out += struct.pack("<d", float(value))
        elif schema == "string":
            b = value.encode("utf-8")
            out += encode_long(len(b)) + b
        elif schema == "bytes":
            out += encode_long(len(value)) + bytes(value)
        return
    t = schema["type"]
    if t == "record":
        for f in schema["fields"]:
            v = value[f["name"]] if f["name"] in value else f.get("default")
            encode(f["type"], v, names, out)
    elif t == "enum":
        out += encode_long(schema["symbols"].index(value))
    elif t == "array":
        if value:
            out += encode_long(len(value))
            for item in value:
                encode(schema["items"], item, names, out)
        out += encode_long(0)
    elif t == "map":
        if value:
            out += encode_long(len(value))
            for k, v in value.items():
                encode("string", k, names, out)
                encode(schema["values"], v, names, out)
        out += encode_long(0)
    else:
        encode(t, value, names, out)


def decode(schema: Any, buf: BinaryIO, names: Names) -> Any:
    if isinstance(schema, list):
        idx = decode_long(buf)
        branch = schema[idx]
        v = decode(branch, buf, names)
        if isinstance(v, dict):
            b = names.resolve(branch) if isinstance(branch, str) and branch not in PRIMITIVES else branch
            if isinstance(b, dict) and b.get("type") == "record":
                v.setdefault("__type__", b["name"])
        return v
    if isinstance(schema, str):
        if schema not in PRIMITIVES:
            return decode(names.resolve(schema), buf, names)
        if schema == "null":
            return None
        if schema == "boolean":
            return _read_exact(buf, 1) != b"\x00"
        if schema in ("int", "long"):
            return decode_long(buf)
        if schema == "float":
            return struct.unpack("<f", _read_exact(buf, 4))[0]
        if schema == "double":
            return struct.unpack("<d", _read_exact(buf, 8))[0]
        if schema == "string":
            return _read_exact(buf, decode_long(buf)).decode("utf-8")
        if schema == "bytes":
            return _read_exact(buf, decode_long(buf))
    t = schema["type"]
    if t == "record":
        return {f["name"]: decode(f["type"], buf, names) for f in schema["fields"]}
    if t == "enum":
        return schema["symbols"][decode_long(buf)]
    if t == "array":
        items: List[Any] = []
        while True:
            n = decode_long(buf)
            if n == 0:
                break
            if n < 0:
                n = -n
                decode_long(buf)  # block byte size
            for _ in range(n):
                items.append(decode(schema["items"], buf, names))
        return items
    if t == "map":
        m: Dict[str, Any] = {}
        while True:
            n = decode_long(buf)
            if n == 0:
                break
            if n < 0:
                n = -n
                decode_long(buf)
            for _ in range(n):
                k = decode("string", buf, names)
                m[k] = decode(schema["values"], buf, names)
        return m
    return decode(t, buf, names)


# ---- object container files ----------------------------------------------------------------
class DataFileWriter:
    """Append-only OCF writer; every ``append`` + ``flush`` leaves a readable file."""

    def __init__(self, path: str, schema: Any, sync: Optional[bytes] = None):
        self.path = path
        self.schema = schema
        self.names = Names()
        _collect(schema, self.names)
        self.sync = sync or os.urandom(SYNC_SIZE)
        self._f = open(path, "wb")
        self._pending: List[bytes] = []
        header = bytearray(MAGIC)
        meta = {"avro.schema": json.dumps(schema, separators=(",", ":")).encode("utf-8")}
        header += encode_long(len(meta))
        for k, v in meta.items():
            kb = k.encode("utf-8")
            header += encode_long(len(kb)) + kb + encode_long(len(v)) + v
        header += encode_long(0)
        header += self.sync
        self._f.write(header)
        self._f.flush()

    def append(self, datum: Any) -> None:
        out = bytearray()
        encode(self.schema, datum, self.names, out)
        self._pending.append(bytes(out))

    def flush(self) -> None:
        if not self._pending:
            return
        body = b"".join(self._pending)
        self._f.write(encode_long(len(self._pending)) + encode_long(len(body)) + body + self.sync)
        self._f.flush()
        os.fsync(self._f.fileno())
        self._pending = []

    def close(self) -> None:
        self.flush()
        self._f.close()


def read_container(path_or_file) -> Tuple[Any, List[Any]]:
    """Return (writer schema, [objects]) of an OCF (null codec only)."""
    f = open(path_or_file, "rb") if isinstance(path_or_file, (str, os.PathLike)) else path_or_file
    try:
        if _read_exact(f, 4) != MAGIC:
            raise ValueError("not an Avro object container file")
        meta: Dict[str, bytes] = {}
        while True:
            n = decode_long(f)
            if n == 0:
                break
            if n < 0:
                n = -n
                decode_long(f)
            for _ in range(n):
                k = _read_exact(f, decode_long(f)).decode("utf-8")
                meta[k] = _read_exact(f, decode_long(f))
        codec = meta.get("avro.codec", b"null").decode()
        if codec != "null":
            raise ValueError(f"unsupported Avro codec {codec!r}")
        schema = json.loads(meta["avro.schema"].decode("utf-8"))
        names = Names()
        _collect(schema, names)
        sync = _read_exact(f, SYNC_SIZE)
        objects: List[Any] = []
        while True:
            head = f.read(1)
            if not head:
                break
            f.seek(-1, io.SEEK_CUR)
            count = decode_long(f)
            size = decode_long(f)
            block = io.BytesIO(_read_exact(f, size))
            for _ in range(count):
                objects.append(decode(schema, block, names))
            if _read_exact(f, SYNC_SIZE) != sync:
                raise ValueError("Avro sync marker mismatch")
        return schema, objects
    finally:
        if isinstance(path_or_file, (str, os.PathLike)):
            f.close()
