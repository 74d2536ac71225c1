"""tcgen05 issue-path micro-probe: cycles per iteration (4 MMAs of 128 x N x 32) for the mode bits of probe_i8_kernel
(1 = commit per iteration, 2 = fence::after_thread_sync, 4 = wait for the commit two iterations back,
 8 = wait for this iteration's commit, 16 = two issuing warps)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tachikoma_b200 import runtime as RT
L = RT.load_library()
L.tkb_probe_int8_shape.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.c_void_p]
torch.cuda.init()
sm, clk = 148, 1.965e9
for n in (64, 128, 256):
    for mode in (0, 1, 3, 16, 17):
        t = ctypes.c_double()
        rc = L.tkb_probe_int8_shape(4000, n, mode, ctypes.byref(t), None)
        if rc != 0:
            print(n, mode, "failed", RT.last_error() if hasattr(RT, "last_error") else "")
            continue
        ops_per_iter = 4 * 2 * 128 * n * 32 * sm * (2 if mode & 16 else 1)
        cyc = ops_per_iter / (t.value * 1e12) * clk
        print(f"N={n:3d} mode={mode:2d}: {t.value:7.0f} TOPS  -> {cyc:6.1f} cycles per 4-MMA iteration per issuing warp ({cyc/4:5.1f} per MMA)")
