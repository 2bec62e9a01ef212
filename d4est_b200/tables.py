"""Reader of flat-table files ("D4BFIX1": name[32] | dtype int32 | pad int32 | count int64 | data), the format
the d4est-side harness writes (oracle/harness/fixture_io.h).  Lets tools load a mesh level without touching
oracle/ (which is test infrastructure)."""
import numpy as np

_DT = {0: np.float64, 1: np.int32, 2: np.int64, 3: np.uint8}
HM_DTYPE = np.dtype([("nbr", "<i4"), ("f_m", "i1"), ("f_p", "i1"), ("tcode", "i1"), ("child_m", "i1"),
                     ("child_p", "i1"), ("degq", "i1"), ("face_slot", "i1"), ("pad", "i1"), ("pad0", "<i4"),
                     ("fac_off", "<i8"), ("g_off", "<i8"), ("pad2", "<i8")])


def load(path):
    out = {}
    with open(path, "rb") as f:
        assert f.read(8) == b"D4BFIX1\0", "not a D4BFIX1 file"
        while True:
            nm = f.read(32)
            if len(nm) < 32:
                break
            dtype, _ = np.frombuffer(f.read(8), dtype=np.int32)
            count = int(np.frombuffer(f.read(8), dtype=np.int64)[0])
            dt = np.dtype(_DT[int(dtype)])
            out[nm.rstrip(b"\0").decode()] = np.frombuffer(f.read(count * dt.itemsize), dtype=dt).copy()
    if "hm" in out:
        out["hm"] = out["hm"].view(HM_DTYPE)
    return out
