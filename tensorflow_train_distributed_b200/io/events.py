"""TensorBoard events files (``tf.summary.scalar`` at distribute_training.py:128-132, written by the chief)."""
from __future__ import annotations

import os
import struct
import time
from typing import Iterator, List, Optional, Tuple

from . import bundle


class EventsWriter:
    def __init__(self, logdir: str, filename_suffix: str = ""):
        os.makedirs(logdir, exist_ok=True)
        self._lib = bundle.lib()
        self._h = self._lib.ttdio_events_open(logdir.encode(), time.time(), filename_suffix.encode())
        if not self._h:
            raise bundle.BundleError("cannot open events file: " + self._lib.ttdio_last_error().decode())
        self.path = self._lib.ttdio_events_path(self._h).decode()

    def add_scalar(self, tag: str, value: float, step: int, wall_time: Optional[float] = None) -> None:
        rc = self._lib.ttdio_events_add_scalar(self._h, tag.encode(), float(value), int(step),
                                               time.time() if wall_time is None else wall_time)
        if rc != 0:
            raise bundle.BundleError(self._lib.ttdio_last_error().decode())

    def flush(self) -> None:
        self._lib.ttdio_events_flush(self._h)

    def close(self) -> None:
        if self._h:
            self._lib.ttdio_events_close(self._h)
            self._h = None


# ---- a small pure-Python reader (tests, and `python -m ... io.events <file>`) -------------------------------
def _varint(buf: bytes, pos: int) -> Tuple[int, int]:
    out = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7


def _fields(buf: bytes) -> Iterator[Tuple[int, int, object]]:
    pos = 0
    while pos < len(buf):
        tag, pos = _varint(buf, pos)
        field, wire = tag >> 3, tag & 7
        if wire == 0:
            v, pos = _varint(buf, pos)
        elif wire == 1:
            v, pos = buf[pos:pos + 8], pos + 8
        elif wire == 2:
            n, pos = _varint(buf, pos)
            v, pos = buf[pos:pos + n], pos + n
        elif wire == 5:
            v, pos = buf[pos:pos + 4], pos + 4
        else:
            raise ValueError(f"unsupported wire type {wire}")
        yield field, wire, v


def read_events(path: str) -> List[dict]:
    """Parse an events file, verifying both CRCs of every record."""
    out = []
    with open(path, "rb") as f:
        data = f.read()
    pos = 0
    while pos < len(data):
        header = data[pos:pos + 8]
        (length,) = struct.unpack("<Q", header)
        (hcrc,) = struct.unpack("<I", data[pos + 8:pos + 12])
        if hcrc != bundle.masked_crc32c(header):
            raise ValueError("events: length checksum mismatch")
        payload = data[pos + 12:pos + 12 + length]
        (dcrc,) = struct.unpack("<I", data[pos + 12 + length:pos + 16 + length])
        if dcrc != bundle.masked_crc32c(payload):
            raise ValueError("events: data checksum mismatch")
        pos += 16 + length
        ev = {"wall_time": 0.0, "step": 0}
        for field, wire, v in _fields(payload):
            if field == 1:
                ev["wall_time"] = struct.unpack("<d", v)[0]
            elif field == 2:
                ev["step"] = v
            elif field == 3:
                ev["file_version"] = v.decode()
            elif field == 5:
                scalars = {}
                for f2, _, val in _fields(v):
                    if f2 == 1:
                        tag, sv = None, None
                        for f3, _, x in _fields(val):
                            if f3 == 1:
                                tag = x.decode()
                            elif f3 == 2:
                                sv = struct.unpack("<f", x)[0]
                        scalars[tag] = sv
                ev["scalars"] = scalars
        out.append(ev)
    return out
