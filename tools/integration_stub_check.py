import ctypes as C, numpy as np, sys
sys.path.insert(0, ".")
L = C.CDLL("rtldavis_b200/lib/libdavis_b200.so")

class DavisCfg(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("bit_rate", "symbol_length", "preamble_symbols",
                "packet_symbols")] + [("preamble", C.c_uint32)] + \
               [(n, C.c_int32) for n in ("block_size", "max_blocks", "max_hits",
                "host_slots", "host_group", "flavor")]    # flavor: 0 Python, 1 Go

HIT = np.dtype([("position", "<i8"), ("stream", "<u4"), ("call", "<u4"), ("index", "<i4"),
                ("data", "u1", (10,)), ("crc_ok", "u1"), ("flags", "u1"),
                ("rssi", "<f8"), ("snr", "<f8"), ("disc_mean", "<f8")])       # 56 bytes

cfg = DavisCfg(19200, 14, 16, 80, 0xCB89, 8192, 1, 0, 0, 0, 0)
h = C.c_void_p()
assert L.davis_create(C.byref(cfg), 1, 0, C.byref(h)) == 0

def demodulate(block_u8):                       # np.uint8[16384], like dsp.py:139
    rc = L.davis_demod_u8(h, block_u8.ctypes.data_as(C.c_void_p), C.c_size_t(block_u8.size),
                          C.c_size_t(block_u8.size), 1, None)
    if rc == -2: raise ValueError("Incompatible array sizes")
    out, n = np.empty(64, HIT), C.c_size_t()
    assert L.davis_fetch_hits(h, out.ctypes.data_as(C.c_void_p), C.c_size_t(64), C.byref(n)) == 0
    return out[:n.value]

g = np.load("tests/golden/hi_snr_8192.npz")
iq = g["iq"]
rows = []
for k in range(iq.size // 16384):
    for x in demodulate(np.ascontiguousarray(iq[k * 16384:(k + 1) * 16384])):
        rows.append((k, int(x["index"]), bytes(x["data"].tolist()), bool(x["crc_ok"])))
want = [(int(c), int(i), bytes(d.tolist()), bool(o)) for c, i, d, o in zip(g["pk_call"], g["pk_index"], g["pk_data"], g["pk_crc"])]
print("stub ok" if rows == want else ("MISMATCH", rows[:3], want[:3]))
try:
    demodulate(np.zeros(100, np.uint8))
except ValueError as e:
    print("size error ok:", e)
