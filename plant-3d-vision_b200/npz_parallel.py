#!/usr/bin/env python3
# -*- coding: utf-8 -*-
"""
plant3dvision_b200.npz_parallel
_______________________________

``savez_compressed_parallel``: the same on-disk format as ``numpy.savez_compressed`` (a ZIP archive of deflated
``<key>.npy`` members, what ``plantdb.io.write_npz`` produces for the ``Voxels`` task, tasks/cl.py:176-180 /
plantdb io.py:610-622) written with all host cores.

numpy deflates each member in one thread (zlib level 6: ~30 MB/s, i.e. minutes for the multi-GB label volumes).  Here
every member is cut into chunks that are deflated concurrently (zlib releases the GIL): each chunk is a raw deflate
stream ended by a sync flush, i.e. a run of non-final blocks closed on a byte boundary, so the concatenation of the
chunks (the last one finished normally) is ONE valid deflate stream, as in pigz.  ``numpy.load`` / ``zipfile`` read the
result like any other ``.npz``; the arrays are bit-identical, only the deflate block boundaries differ.
"""
import io
import os
import struct
import time
import zlib
from concurrent.futures import ThreadPoolExecutor

import numpy as np

CHUNK = 4 << 20
_MAX32 = 0xFFFFFFFF


def _npy_header(arr):
    b = io.BytesIO()
    d = np.lib.format.header_data_from_array_1_0(arr)
    try:
        np.lib.format.write_array_header_1_0(b, d)
    except ValueError:  # header too long for version 1.0
        b = io.BytesIO()
        np.lib.format.write_array_header_2_0(b, d)
    return b.getvalue()


def _deflate(chunk, level, last):
    co = zlib.compressobj(level, zlib.DEFLATED, -15)
    out = co.compress(chunk)
    return out + co.flush(zlib.Z_FINISH if last else zlib.Z_SYNC_FLUSH)


def _dos_time(t=None):
    lt = time.localtime(t)
    year = max(lt.tm_year, 1980)
    return (lt.tm_hour << 11) | (lt.tm_min << 5) | (lt.tm_sec // 2), ((year - 1980) << 9) | (lt.tm_mon << 5) | lt.tm_mday


def savez_compressed_parallel(file, arrays, level=6, threads=None, chunk=CHUNK, force_zip64=False, window=None):
    """Write ``arrays`` (mapping key -> ndarray) to ``file`` (a path) as ``numpy.savez_compressed`` would.
    ZIP64 records are written when a size or offset needs them (``force_zip64``: always, as numpy does).
    At most ``window`` chunks (default 4 per thread) are in flight, so memory stays bounded whatever the volume
    size: a member's local header is written first and patched with its CRC and sizes once its last chunk is out."""
    threads = threads or min(32, os.cpu_count() or 1)
    window = max(1, window or 4 * threads)
    path = os.fspath(file)
    if not path.endswith(".npz"):
        path += ".npz"  # numpy.savez appends the extension too
    tm, dt = _dos_time()
    central = []
    with ThreadPoolExecutor(max_workers=threads) as pool, open(path, "wb") as f:
        for key, arr in arrays.items():
            arr = np.asanyarray(arr)
            if arr.dtype.hasobject:
                raise TypeError("object arrays are not supported")
            if not arr.flags["C_CONTIGUOUS"]:
                arr = np.ascontiguousarray(arr)  # numpy writes Fortran arrays as such; C order keeps np.load equal
            head = _npy_header(arr)
            data = memoryview(arr.reshape(-1)).cast("B") if arr.size else memoryview(b"")
            n = len(data)
            usize = len(head) + n
            n_pieces = 1 + (n + chunk - 1) // chunk

            def piece(i, head=head, data=data):
                return head if i == 0 else data[(i - 1) * chunk:i * chunk]

            nameb = (key + ".npy").encode("utf-8")
            flags = 0x800 if any(ord(ch) > 127 for ch in key) else 0
            # deflate never grows data by more than a few bytes per 16 KB block: sizes need ZIP64 only near 4 GiB
            z64 = force_zip64 or usize >= _MAX32 - (1 << 24)
            offset = f.tell()

            def local_header(crc, csize):
                extra = struct.pack("<HHQQ", 1, 16, usize, csize) if z64 else b""
                return struct.pack("<IHHHHHIIIHH", 0x04034B50, 45 if z64 else 20, flags, 8, tm, dt, crc,
                                   _MAX32 if z64 else csize, _MAX32 if z64 else usize, len(nameb), len(extra)) + nameb + extra

            f.write(local_header(0, 0))
            crc, csize, pending, nxt = 0, 0, [], 0
            while nxt < n_pieces or pending:
                while nxt < n_pieces and len(pending) < window:
                    pending.append(pool.submit(_deflate, piece(nxt), level, nxt == n_pieces - 1))
                    crc = zlib.crc32(piece(nxt), crc)  # runs beside the deflate workers
                    nxt += 1
                comp = pending.pop(0).result()
                f.write(comp)
                csize += len(comp)
            end = f.tell()
            f.seek(offset)
            f.write(local_header(crc, csize))
            f.seek(end)
            central.append((nameb, flags, crc, csize, usize, offset, z64))
        cd_off = f.tell()
        for nameb, flags, crc, csize, usize, offset, z64sizes in central:
            ex = b""
            fields = []
            z64off = force_zip64 or offset >= _MAX32
            if z64sizes:
                fields += [usize, csize]
            if z64off:
                fields.append(offset)
            if fields:
                ex = struct.pack("<HH" + "Q" * len(fields), 1, 8 * len(fields), *fields)
            f.write(struct.pack("<IHHHHHHIIIHHHHHII", 0x02014B50, (3 << 8) | 45, 45 if fields else 20, flags, 8, tm,
                                dt, crc, _MAX32 if z64sizes else csize, _MAX32 if z64sizes else usize, len(nameb),
                                len(ex), 0, 0, 0, 0o600 << 16, _MAX32 if z64off else offset))
            f.write(nameb)
            f.write(ex)
        cd_size = f.tell() - cd_off
        n_ent = len(central)
        # (an empty archive stays a bare end-of-central-directory record: numpy recognises it by its magic)
        big = (force_zip64 and n_ent > 0) or n_ent >= 0xFFFF or cd_off >= _MAX32 or cd_size >= _MAX32
        if big:
            z_off = f.tell()
            f.write(struct.pack("<IQHHIIQQQQ", 0x06064B50, 44, 45, 45, 0, 0, n_ent, n_ent, cd_size, cd_off))
            f.write(struct.pack("<IIQI", 0x07064B50, 0, z_off, 1))
        f.write(struct.pack("<IHHHHIIH", 0x06054B50, 0, 0, 0xFFFF if big else n_ent, 0xFFFF if big else n_ent,
                            _MAX32 if big else cd_size, _MAX32 if big else cd_off, 0))
    return path
