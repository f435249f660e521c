import sys; sys.path.insert(0,'.')
import numpy as np
from rtldavis_b200 import dsp, protocol, synth, _lib
B=8192
cfg=protocol.new_packet_config(14)
iq,_=synth.make_batch(4,16*B,8,30.0,seed0=8200)
for mh in (8, 0):
    bd=dsp.BatchDemodulator(cfg,4,max_blocks=8,max_hits=mh)
    bd.submit(iq[:, :2*B*8])
    try:
        h=bd.fetch(); print("first ok", len(h))
    except Exception as e: print("first:", str(e)[:90])
    st=bd.stats(); print(" stats", st.candidates, st.confirmed, st.records, st.reruns)
    for k in range(8,12):
        bd.submit(iq[:, 2*B*k:2*B*(k+1)])
        try:
            h=bd.fetch(); print(k, "ok", len(h))
        except Exception as e: print(k, str(e)[:120])
        st=bd.stats(); print("   stats", st.candidates, st.confirmed, st.records, st.reruns)
    bd.close()
