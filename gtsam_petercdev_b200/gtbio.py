"""GTB1 named-array container (little endian), shared with oracle/gtb_io.h.

magic "GTB1", u32 n_records, per record: u32 name_len, name, u8 dtype
(0=f64 1=i64 2=i32 3=u64), u32 ndim, u64 shape[ndim], raw C-order data.
"""
import struct
import numpy as np

_DT = {0: np.float64, 1: np.int64, 2: np.int32, 3: np.uint64}
_CODE = {np.dtype(v): k for k, v in _DT.items()}


def save(path, arrays):
    with open(path, "wb") as f:
        f.write(b"GTB1")
        f.write(struct.pack("<I", len(arrays)))
        for name, a in arrays.items():
            a = np.ascontiguousarray(a)
            if a.dtype not in _CODE:
                raise TypeError(f"{name}: unsupported dtype {a.dtype}")
            nb = name.encode()
            f.write(struct.pack("<I", len(nb)))
            f.write(nb)
            f.write(struct.pack("<BI", _CODE[a.dtype], a.ndim))
            f.write(struct.pack(f"<{a.ndim}Q", *a.shape))
            f.write(a.tobytes())


def load(path):
    out = {}
    with open(path, "rb") as f:
        if f.read(4) != b"GTB1":
            raise ValueError("bad magic")
        (n,) = struct.unpack("<I", f.read(4))
        for _ in range(n):
            (ln,) = struct.unpack("<I", f.read(4))
            name = f.read(ln).decode()
            code, nd = struct.unpack("<BI", f.read(5))
            shape = struct.unpack(f"<{nd}Q", f.read(8 * nd))
            dt = np.dtype(_DT[code])
            cnt = int(np.prod(shape)) if nd else 1
            out[name] = np.frombuffer(f.read(cnt * dt.itemsize), dtype=dt).reshape(shape).copy()
    return out
