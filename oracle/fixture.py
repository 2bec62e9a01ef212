"""Reader/writer for oracle fixture files (TEST INFRASTRUCTURE ONLY).

Format written by oracle/harness/fixture_io.h: "D4BFIX1\\0" then records
name[32] | dtype int32 | pad int32 | count int64 | data.
"""
import numpy as np

_DT = {0: np.float64, 1: np.int32, 2: np.int64, 3: np.uint8}
_CODE = {np.dtype(v): k for k, v in _DT.items()}

# numpy view of d4b_hm_t (include/d4est_b200.h), 40 bytes
HM_DTYPE = np.dtype([("nbr", "<i4"), ("f_m", "i1"), ("f_p", "i1"), ("tcode", "i1"),
                     ("child_m", "i1"), ("child_p", "i1"), ("degq", "i1"),
                     ("face_slot", "i1"), ("pad", "i1"), ("pad0", "<i4"), ("fac_off", "<i8"),
                     ("g_off", "<i8"), ("pad2", "<i8")])


def load(path):
    out = {}
    with open(path, "rb") as f:
        assert f.read(8) == b"D4BFIX1\0", "not a D4BFIX1 file"
        while True:
            nm = f.read(32)
            if len(nm) < 32:
                break
            dtype, _pad = np.frombuffer(f.read(8), dtype=np.int32)
            count = int(np.frombuffer(f.read(8), dtype=np.int64)[0])
            dt = np.dtype(_DT[int(dtype)])
            out[nm.rstrip(b"\0").decode()] = np.frombuffer(f.read(count * dt.itemsize), dtype=dt).copy()
    for k in list(out):
        if k == "hm" or k.endswith(".hm"):
            out[k] = out[k].view(HM_DTYPE)
    return out


def sub(fx, prefix):
    """The arrays of one level of a multigrid fixture ("L2." ...) as a plain level fixture."""
    return {k[len(prefix):]: v for k, v in fx.items() if k.startswith(prefix)}


def save(path, arrays):
    with open(path, "wb") as f:
        f.write(b"D4BFIX1\0")
        for name, a in arrays.items():
            a = np.ascontiguousarray(a)
            if a.dtype == HM_DTYPE:
                a = a.view(np.uint8)
            nm = name.encode()[:31].ljust(32, b"\0")
            f.write(nm)
            f.write(np.array([_CODE[a.dtype], 0], dtype=np.int32).tobytes())
            f.write(np.array([a.size], dtype=np.int64).tobytes())
            f.write(a.tobytes())
