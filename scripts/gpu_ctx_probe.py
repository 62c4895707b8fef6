"""CUDA start-up cost as the CLI sees it: library load, first context, second context."""
import sys, time
sys.path.insert(0, ".")
t0 = time.perf_counter()
import numpy as np
t1 = time.perf_counter()
from reportree_b200 import hc_backend as hc
L = hc.load_library()
t2 = time.perf_counter()
ctx = hc.HCContext(device=0)
t3 = time.perf_counter()
ctx.close()
t4 = time.perf_counter()
ctx = hc.HCContext(device=0)
t5 = time.perf_counter()
print("numpy import %.3f s, library load %.3f s, first context %.3f s, close %.3f s, second context %.3f s"
      % (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4))
