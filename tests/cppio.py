"""Named-array container shared with tests/cpp/blobio.h (C++ drop-in tests)."""
import numpy as np

_DT = {"u1": np.uint8, "i4": np.int32, "f4": np.float32, "u4": np.uint32}


def write(path, arrays):
    with open(path, "wb") as f:
        for name, a in arrays.items():
            a = np.ascontiguousarray(a)
            code = {np.dtype(np.uint8): "u1", np.dtype(np.int32): "i4", np.dtype(np.float32): "f4", np.dtype(np.uint32): "u4"}[a.dtype]
            f.write(("%s %s %d\n" % (name, code, a.size)).encode())
            f.write(a.tobytes())


def read(path):
    out = {}
    with open(path, "rb") as f:
        while True:
            line = f.readline()
            if not line:
                break
            name, code, n = line.decode().split()
            dt = np.dtype(_DT[code])
            out[name] = np.frombuffer(f.read(int(n) * dt.itemsize), dt).copy()
    return out
