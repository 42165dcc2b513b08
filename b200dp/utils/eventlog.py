"""Rank-0 scalar event log, readable by TensorBoard.

The reference attaches ``tf.keras.callbacks.TensorBoard(log_dir='./logs')`` on rank
0 only (``horovod/tensorflow_mnist_gpu.py:156-158``).  That is plain file I/O: a
TFRecord stream of ``Event`` protos.  This writer emits the same on-disk format
(hand-encoded protobuf + masked CRC32C framing — no TensorFlow needed) and a
``scalars.jsonl`` twin for tools that prefer JSON.
"""
from __future__ import annotations

import json
import os
import socket
import struct
import time
from typing import Dict, Optional

_CRC_TABLE = []


def _crc_table():
    if not _CRC_TABLE:
        poly = 0x82F63B78
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ poly if c & 1 else c >> 1
            _CRC_TABLE.append(c)
    return _CRC_TABLE


def crc32c(data: bytes) -> int:
    t = _crc_table()
    c = 0xFFFFFFFF
    for b in data:
        c = t[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def _masked(data: bytes) -> int:
    c = crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


def _varint(n: int) -> bytes:
    out = bytearray()
    n &= (1 << 64) - 1
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _field_bytes(num: int, payload: bytes) -> bytes:
    return _varint((num << 3) | 2) + _varint(len(payload)) + payload


def encode_event(wall_time: float, step: int = 0, file_version: Optional[str] = None,
                 scalars: Optional[Dict[str, float]] = None) -> bytes:
    ev = _varint((1 << 3) | 1) + struct.pack("<d", wall_time)
    if step:
        ev += _varint((2 << 3) | 0) + _varint(step)
    if file_version is not None:
        ev += _field_bytes(3, file_version.encode())
    if scalars:
        summ = b""
        for tag, val in scalars.items():
            v = _field_bytes(1, tag.encode()) + _varint((2 << 3) | 5) + struct.pack("<f", float(val))
            summ += _field_bytes(1, v)
        ev += _field_bytes(5, summ)
    return ev


def frame_record(data: bytes) -> bytes:
    hdr = struct.pack("<Q", len(data))
    return hdr + struct.pack("<I", _masked(hdr)) + data + struct.pack("<I", _masked(data))


class EventWriter:
    def __init__(self, log_dir: str):
        os.makedirs(log_dir, exist_ok=True)
        name = f"events.out.tfevents.{int(time.time())}.{socket.gethostname()}.{os.getpid()}.b200dp"
        self.path = os.path.join(log_dir, name)
        self._f = open(self.path, "ab")
        self._j = open(os.path.join(log_dir, "scalars.jsonl"), "a", buffering=1)
        self._f.write(frame_record(encode_event(time.time(), file_version="brain.Event:2")))
        self._f.flush()

    def add_scalars(self, scalars: Dict[str, float], step: int) -> None:
        now = time.time()
        self._f.write(frame_record(encode_event(now, step=step, scalars=scalars)))
        self._f.flush()
        self._j.write(json.dumps({"wall_time": now, "step": step, **{k: float(v) for k, v in scalars.items()}}) + "\n")

    def close(self) -> None:
        for f in (self._f, self._j):
            try:
                f.close()
            except Exception:
                pass


def read_events(path: str):
    """Parse a file written by :class:`EventWriter` back into ``(step, {tag: value})``
    tuples, verifying both CRCs (used by the tests)."""
    out = []
    with open(path, "rb") as f:
        buf = f.read()
    pos = 0
    while pos < len(buf):
        (n,) = struct.unpack_from("<Q", buf, pos)
        (hc,) = struct.unpack_from("<I", buf, pos + 8)
        assert hc == _masked(buf[pos:pos + 8]), "length CRC mismatch"
        data = buf[pos + 12:pos + 12 + n]
        (dc,) = struct.unpack_from("<I", buf, pos + 12 + n)
        assert dc == _masked(data), "data CRC mismatch"
        pos += 16 + n
        out.append(_decode_event(data))
    return out


def _read_varint(b: bytes, i: int):
    n = shift = 0
    while True:
        c = b[i]
        i += 1
        n |= (c & 0x7F) << shift
        shift += 7
        if not c & 0x80:
            return n, i


def _decode_event(b: bytes):
    i = 0
    step = 0
    scalars: Dict[str, float] = {}
    while i < len(b):
        key, i = _read_varint(b, i)
        num, wt = key >> 3, key & 7
        if wt == 1:
            i += 8
        elif wt == 0:
            v, i = _read_varint(b, i)
            if num == 2:
                step = v
        elif wt == 5:
            i += 4
        elif wt == 2:
            ln, i = _read_varint(b, i)
            payload = b[i:i + ln]
            i += ln
            if num == 5:
                j = 0
                while j < len(payload):
                    k2, j = _read_varint(payload, j)
                    l2, j = _read_varint(payload, j)
                    val = payload[j:j + l2]
                    j += l2
                    q = 0
                    tag, sv = None, None
                    while q < len(val):
                        k3, q = _read_varint(val, q)
                        if k3 & 7 == 2:
                            l3, q = _read_varint(val, q)
                            if k3 >> 3 == 1:
                                tag = val[q:q + l3].decode()
                            q += l3
                        elif k3 & 7 == 5:
                            if k3 >> 3 == 2:
                                (sv,) = struct.unpack_from("<f", val, q)
                            q += 4
                        else:
                            _, q = _read_varint(val, q)
                    if tag is not None:
                        scalars[tag] = sv
    return step, scalars
