import sys, os, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, vulcan_b200 as vb
vb.lib.vk_div_last_check.argtypes=[C.POINTER(C.c_uint*2)]
vb.lib.vk_div_last_check.restype=None
for d in (0.004, 0.032, 0.02, 0.008, 640.0, 480.0, 3.0, 0.0123456):
    ok = vb.lib.vk_div_prepare(np.float32(d), None)
    out=(C.c_uint*2)(); vb.lib.vk_div_last_check(C.byref(out))
    x=np.array([out[1]],dtype=np.uint32).view(np.float32)[0]
    print(d, ok, out[0], hex(out[1]), x)
