#!/bin/bash
python - <<'PY'
import ctypes
from tachikoma_b200 import runtime as RT
L = RT.load_library()
L.tkb_probe_int8_shape.argtypes=[ctypes.c_int,ctypes.c_int,ctypes.c_int,ctypes.POINTER(ctypes.c_double),ctypes.c_void_p]
for n in (256,128,64):
    for ce in (0,1,2,3,5,7,9):
        t=ctypes.c_double()
        rc=L.tkb_probe_int8_shape(4000,n,ce,ctypes.byref(t),None)
        per_mma_ns = (2*128*n*32*148)/(t.value*1e12)*1e9 if t.value>0 else -1
        print(f"N={n} mode={ce}: {t.value:.0f} TOPS  -> {4*per_mma_ns*1.965:.0f} cycles per 4-MMA iteration (rc={rc})")
PY
