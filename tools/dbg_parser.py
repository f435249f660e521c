import sys; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import numpy as np, logging
logging.disable(logging.CRITICAL)
from conftest import load_seeded
from oracle import ref_loader
from rtldavis_b200 import dsp as gd, protocol as gp
_,_,rp = ref_loader.load()
g, iq = load_seeded("config1_10s")
B=8192
pr = rp.Parser(symbol_length=14)
dem = gd.Demodulator(gp.new_packet_config(14))
for k in range(12):
    blk = iq[2*B*k:2*B*(k+1)]
    a = pr.demodulator.demodulate(blk)
    b = dem.demodulate(blk)
    if a:
        d1 = pr.demodulator.discriminated; d2 = dem.discriminated
        i = a[0].index
        print(k, i, b[0].index, float(np.mean(d1[i:i+224])), float(np.mean(d2[i:i+224])), np.abs(d1-d2).max(), d2[:3], d2[-3:])
