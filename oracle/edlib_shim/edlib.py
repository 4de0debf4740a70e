"""Stand-in for python-edlib so the UNMODIFIED reference (py/sequence.py, py/tksm_badread.py) runs here.

TEST INFRASTRUCTURE.  Put this directory on PYTHONPATH ahead of /root/reference/py.  Only
`align(query, target, mode="NW", task="path")` is provided, which is all the reference uses
(py/tksm_badread.py:170, 409-411, 422-426, 613).  The alignment itself is oracle/nw.c.
"""
import ctypes
import os
import subprocess

_here = os.path.dirname(os.path.abspath(__file__))
_so = os.path.join(_here, "..", "_build", "libnw.so")
if not os.path.exists(_so):
    subprocess.check_call(["make", "-C", os.path.join(_here, ".."), "_build/libnw.so"])
_lib = ctypes.CDLL(_so)
_lib.nw_align.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p,
                          ctypes.POINTER(ctypes.c_int)]
_lib.nw_align.restype = ctypes.c_int


def align(query, target, mode="NW", task="distance", k=-1, additionalEqualities=None):
    assert mode == "NW"
    q = query.encode("latin-1") if isinstance(query, str) else bytes(query)
    t = target.encode("latin-1") if isinstance(target, str) else bytes(target)
    buf = ctypes.create_string_buffer(len(q) + len(t) + 1)
    dist = ctypes.c_int(0)
    n = _lib.nw_align(q, len(q), t, len(t), buf, ctypes.byref(dist))
    if n < 0:
        raise MemoryError("nw_align")
    ops = buf.raw[:n].decode()
    cigar = []
    i = 0
    while i < n:
        j = i
        while j < n and ops[j] == ops[i]:
            j += 1
        cigar.append(f"{j - i}{ops[i]}")
        i = j
    return {"editDistance": dist.value, "cigar": "".join(cigar) if task == "path" else None,
            "alphabetLength": 4, "locations": [(None, len(t) - 1)]}
