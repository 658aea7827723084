import time, os, sys
sys.path.insert(0, os.getcwd())
t0=time.perf_counter()
import numpy as np
from ecoz2rs_b200 import vq
t1=time.perf_counter()
L=vq.lib(); n=vq.device_count()
t2=time.perf_counter()
from ecoz2rs_b200.synth import ar_mix
x=ar_mix(100000,36,seed=1)
t3=time.perf_counter()
eng=vq.VqEngine(36,max_M=256)
t4=time.perf_counter()
eng.set_frames(x)
t5=time.perf_counter()
eng.codebook_initial(); out=eng.kmeans_pass()
t6=time.perf_counter()
st,_=eng.learn(eps=0.05, stop_M=256)
t7=time.perf_counter()
eng.codebook_initial(); st,_=eng.learn(eps=0.05, stop_M=256)
t8=time.perf_counter()
print(f"import {t1-t0:.3f} lib+device_count {t2-t1:.3f} engine create {t4-t3:.3f} set_frames {t5-t4:.3f} first pass {t6-t5:.3f} first learn {t7-t6:.3f} second learn {t8-t7:.3f}")
