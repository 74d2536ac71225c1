#!/bin/bash
python - <<'PY'
import ctypes
from tachikoma_b200 import runtime as RT
L = RT.load_library()
L.tkb_probe_int8_shape.argtypes=[ctypes.c_int,ctypes.c_int,ctypes.c_int,ctypes.POINTER(ctypes.c_double),ctypes.c_void_p]
for n in (256,128,64,32):
    for ce in (0,16):
        t=ctypes.c_double()
        rc=L.tkb_probe_int8_shape(4000,n,ce,ctypes.byref(t),None)
        print(f"N={n} mode={ce}: {t.value:.0f} TOPS (rc={rc})")
PY
