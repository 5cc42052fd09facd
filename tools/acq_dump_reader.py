#!/usr/bin/env python3
"""What the reference's Acquisition_Dump_Reader does with an acquisition dump, restated without matio or HDF5.

The reference reads its dumps with matio (src/tests/unit-tests/signal-processing-blocks/libs/acquisition_dump_reader.cc:26-128):
Mat_Open() detects the container from the 128-byte header (matio reads Level 4, Level 5 and 7.3 files through the same calls),
Mat_VarRead(name) walks the data elements until the array name matches, and read_binary_acq() then
  * rejects a grid whose rank is not 2, whose dims are not [samples_per_code, num_doppler_bins], or whose data type is not single;
  * takes doppler_max / doppler_step through an `unsigned int*`, input_power / acq_doppler_hz / acq_delay_samples / test_statistic /
    threshold through a `float*`, sample_counter through a `uint64_t*`, d_positive_acq / num_dwells / PRN through an `int*`;
  * fills mag[doppler][sample] = grid[k] / (samples_per_code^4 * input_power), k running over the column-major data.
matio is not installed in this image, so this module parses the Level-5 container itself, from the published "MAT-File Format"
document (header, miMATRIX elements: array flags, dimensions, name, real part; small-element form; 8-byte padding; miCOMPRESSED
elements through zlib), and applies exactly the checks and pointer casts above. It is a check that the files
host/b200_mat_writer produces carry what that reader takes, in the classes it takes them — not a run of matio itself.
"""
import struct
import sys
import zlib

import numpy as np

MI = {1: ("i1", 1), 2: ("u1", 1), 3: ("i2", 2), 4: ("u2", 2), 5: ("i4", 4), 6: ("u4", 4), 7: ("f4", 4), 9: ("f8", 8), 12: ("i8", 8), 13: ("u8", 8)}
MI_MATRIX, MI_COMPRESSED = 14, 15
# mxCLASS -> the matio data type of a numeric array of that class
MX = {6: "f8", 7: "f4", 8: "i1", 9: "u1", 10: "i2", 11: "u2", 12: "i4", 13: "u4", 14: "i8", 15: "u8"}


class MatError(Exception):
    pass


def _tag(buf, pos, e):
    """(type, nbytes, payload offset, offset of the next element)"""
    word, = struct.unpack_from(e + "I", buf, pos)
    if word >> 16:  # small data element: 2-byte size, 2-byte type, 4 bytes of payload
        return word & 0xFFFF, word >> 16, pos + 4, pos + 8
    t, n = struct.unpack_from(e + "II", buf, pos)
    nxt = pos + 8 + n
    if t != MI_COMPRESSED:
        nxt = (nxt + 7) & ~7
    return t, n, pos + 8, nxt


def _matrix(buf, e):
    pos = 0
    t, n, off, pos = _tag(buf, pos, e)
    if t != 6 or n != 8:
        raise MatError("array flags missing")
    flags, = struct.unpack_from(e + "I", buf, off)
    cls, is_complex = flags & 0xFF, bool(flags & 0x0800)
    t, n, off, pos = _tag(buf, pos, e)
    if t != 5:
        raise MatError("dimensions array missing")
    dims = struct.unpack_from(e + "%di" % (n // 4), buf, off)
    t, n, off, pos = _tag(buf, pos, e)
    if t != 1:
        raise MatError("array name missing")
    name = bytes(buf[off:off + n]).decode("ascii")
    if cls not in MX or is_complex:
        return name, None
    t, n, off, pos = _tag(buf, pos, e)
    if t not in MI:
        raise MatError("unknown data type %d in %s" % (t, name))
    stored = np.frombuffer(buf, dtype=np.dtype(e + MI[t][0]), count=n // MI[t][1], offset=off)
    # matio converts the stored element type to the class type on read (Mat_VarRead -> ReadData)
    data = stored.astype(np.dtype(MX[cls]))
    return name, dict(rank=len(dims), dims=dims, data_type=MX[cls], data=data)


def mat_open(path):
    """Mat_Open + a walk over every element: {name: matvar}."""
    raw = open(path, "rb").read()
    if len(raw) < 128:
        raise MatError("not a MAT file")
    ver_le, = struct.unpack_from("<H", raw, 124)
    if raw[126:128] == b"IM" and ver_le == 0x0100:
        e = "<"
    elif raw[126:128] == b"MI" and struct.unpack_from(">H", raw, 124)[0] == 0x0100:
        e = ">"
    else:
        raise MatError("not a Level-5 MAT file (version %#x, endian %r)" % (ver_le, raw[126:128]))
    out, pos = {}, 128
    while pos + 8 <= len(raw):
        t, n, off, pos = _tag(raw, pos, e)
        body = raw[off:off + n]
        if t == MI_COMPRESSED:
            inner = zlib.decompress(body)
            t2, n2, off2, _ = _tag(inner, 0, e)
            if t2 != MI_MATRIX:
                continue
            body = inner[off2:off2 + n2]
        elif t != MI_MATRIX:
            continue
        name, var = _matrix(memoryview(body), e)
        if var is not None:
            out[name] = var
    return out


def read_binary_acq(path, samples_per_code, num_doppler_bins):
    """Acquisition_Dump_Reader::read_binary_acq(); raises MatError where the reader returns false."""
    m = mat_open(path)
    if "acq_grid" not in m:
        raise MatError("unreachable grid variable")
    g = m["acq_grid"]
    if g["rank"] != 2:
        raise MatError("rank error")
    if g["dims"][0] != samples_per_code or g["dims"][1] != num_doppler_bins:
        raise MatError("dimension matrix error: %r" % (g["dims"],))
    if g["data_type"] != "f4":
        raise MatError("data type error")

    def through(name, ctype):  # *static_cast<T*>(var->data)
        if name not in m:
            raise MatError("variable %s missing (the reader would dereference a null pointer)" % name)
        b = m[name]["data"].tobytes()
        n = np.dtype(ctype).itemsize
        if len(b) < n:
            raise MatError("%s holds %d bytes, the reader takes %d" % (name, len(b), n))
        return np.frombuffer(b[:n], dtype=ctype)[0]

    out = dict(doppler_max=int(through("doppler_max", "u4")), doppler_step=int(through("doppler_step", "u4")),
               input_power=float(through("input_power", "f4")), acq_doppler_hz=float(through("acq_doppler_hz", "f4")),
               acq_delay_samples=float(through("acq_delay_samples", "f4")), test_statistic=float(through("test_statistic", "f4")),
               threshold=float(through("threshold", "f4")), sample_counter=int(through("sample_counter", "u8")),
               positive_acq=int(through("d_positive_acq", "i4")), num_dwells=int(through("num_dwells", "i4")), PRN=int(through("PRN", "i4")))
    norm = np.float32(np.float32(samples_per_code) ** 4 * np.float32(out["input_power"]))
    out["mag"] = (g["data"].reshape(num_doppler_bins, samples_per_code) / norm).astype(np.float32)   # mag[doppler][sample]
    return out


if __name__ == "__main__":
    r = read_binary_acq(sys.argv[1], int(sys.argv[2]), int(sys.argv[3]))
    print({k: v for k, v in r.items() if k != "mag"}, r["mag"].shape, float(r["mag"].max()))
