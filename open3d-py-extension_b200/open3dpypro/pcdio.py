"""Minimal PCD reader (ascii / binary, uncompressed) so that config 1 (`data/bunny.pcd`, fields
`Intensity x y z _`, reference PointCloud.py:165-166) loads without Open3D."""
import numpy as np

_NP = {("F", 4): "f4", ("F", 8): "f8", ("U", 1): "u1", ("U", 2): "u2", ("U", 4): "u4", ("U", 8): "u8",
       ("I", 1): "i1", ("I", 2): "i2", ("I", 4): "i4", ("I", 8): "i8"}


def read_pcd(path):
    """-> (xyz float64 (n,3), intensity float64 (n,) or None)"""
    with open(path, "rb") as f:
        raw = f.read()
    hdr, pos = {}, 0
    while True:
        end = raw.index(b"\n", pos)
        line = raw[pos:end].decode("ascii", "replace").strip()
        pos = end + 1
        if not line or line.startswith("#"):
            continue
        key, _, val = line.partition(" ")
        hdr[key.upper()] = val.split()
        if key.upper() == "DATA":
            break
    fields = hdr["FIELDS"]
    sizes = [int(v) for v in hdr["SIZE"]]
    types = hdr["TYPE"]
    counts = [int(v) for v in hdr.get("COUNT", ["1"] * len(fields))]
    n = int(hdr["POINTS"][0]) if "POINTS" in hdr else int(hdr["WIDTH"][0]) * int(hdr["HEIGHT"][0])
    mode = hdr["DATA"][0].lower()
    if mode == "binary":
        dt = np.dtype([(name if name != "_" else f"_pad{i}", _NP[(t, s)], (c,) if c > 1 else ())
                       for i, (name, s, t, c) in enumerate(zip(fields, sizes, types, counts))])
        rec = np.frombuffer(raw, dtype=dt, count=n, offset=pos)
        col = lambda k: rec[k].astype(np.float64)
    elif mode == "ascii":
        tab = np.loadtxt(raw[pos:].decode("ascii").splitlines(), ndmin=2)[:n]
        offs = np.cumsum([0] + counts)
        col = lambda k: tab[:, offs[fields.index(k)]].astype(np.float64)
    else:
        raise ValueError(f"unsupported PCD DATA mode {mode!r}")
    low = {f.lower(): f for f in fields}
    xyz = np.stack([col(low["x"]), col(low["y"]), col(low["z"])], 1)
    inten = col(low["intensity"]) if "intensity" in low else None
    return np.ascontiguousarray(xyz), inten
