"""ctypes binding of libtkb200.so -- plays the role the GraphExecutor / ConstLoaderModule play for the
reference's TachikomaJSONRuntime (src/runtime/graph_executor/graph_executor.cc:524-562,
src/runtime/const_loader_module.cc:124-131): create the module from (symbol, graph_json,
const_names), bind the constants once through ``__init_<symbol>``, then call ``<symbol>`` with the
inputs followed by caller-allocated outputs.

There is no CPU fallback: if the CUDA library is missing or no CUDA device is present the calls
raise.  PyTorch is used only to own device memory / streams in the convenience helpers."""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

kDLCPU, kDLCUDA, kDLCUDAHost = 1, 2, 3


class DLDevice(ctypes.Structure):
    _fields_ = [("device_type", ctypes.c_int32), ("device_id", ctypes.c_int32)]


class DLDataType(ctypes.Structure):
    _fields_ = [("code", ctypes.c_uint8), ("bits", ctypes.c_uint8), ("lanes", ctypes.c_uint16)]


class DLTensor(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("device", DLDevice), ("ndim", ctypes.c_int32), ("dtype", DLDataType),
                ("shape", ctypes.POINTER(ctypes.c_int64)), ("strides", ctypes.POINTER(ctypes.c_int64)),
                ("byte_offset", ctypes.c_uint64)]


_DT = {"int8": (0, 8), "uint8": (1, 8), "int16": (0, 16), "int32": (0, 32), "int64": (0, 64), "float32": (2, 32)}


class TachikomaError(RuntimeError):
    """Raised with the message of tkb_last_error(), the analogue of tvm.error.TVMError."""


def lib_path():
    return os.path.join(_HERE, "libtkb200.so")


def load_library():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise TachikomaError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(there is no CPU fallback for the tachikoma CUDA runtime)")
    L = ctypes.CDLL(path)
    H = ctypes.c_void_p
    L.tkb_last_error.restype = ctypes.c_char_p
    L.tkb_type_key.restype = ctypes.c_char_p
    L.tkb_version.restype = ctypes.c_char_p
    L.tkb_get_symbol.restype = ctypes.c_char_p
    L.tkb_get_symbol.argtypes = [H]
    L.tkb_plan_summary.restype = ctypes.c_char_p
    L.tkb_plan_summary.argtypes = [H]
    L.tkb_create.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_char_p), ctypes.c_int,
                             ctypes.POINTER(H)]
    L.tkb_free.argtypes = [H]
    L.tkb_save_binary.argtypes = [H, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_size_t)]
    L.tkb_load_binary.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(H)]
    L.tkb_get_const_vars.argtypes = [H, ctypes.POINTER(ctypes.POINTER(ctypes.c_char_p)), ctypes.POINTER(ctypes.c_int)]
    L.tkb_num_inputs.argtypes = [H]
    L.tkb_num_outputs.argtypes = [H]
    L.tkb_init.argtypes = [H, ctypes.POINTER(ctypes.POINTER(DLTensor)), ctypes.c_int]
    L.tkb_run.argtypes = [H, ctypes.POINTER(ctypes.POINTER(DLTensor)), ctypes.c_int, ctypes.c_void_p]
    L.tkb_run_traced.argtypes = [H, ctypes.POINTER(ctypes.POINTER(DLTensor)), ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    L.tkb_trace_sync_on.argtypes = [H, ctypes.c_void_p]
    L.tkb_trace_blob_on.argtypes = [H, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_size_t)]
    L.tkb_trace_dump_debug.argtypes = [H, ctypes.c_char_p]
    L.tkb_packed_run.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.c_int, ctypes.c_void_p,
                                 ctypes.POINTER(ctypes.c_int), ctypes.c_void_p]
    L.tkb_set_option.argtypes = [H, ctypes.c_char_p, ctypes.c_char_p]
    L.tkb_launch_count.restype = ctypes.c_int64
    L.tkb_launch_count.argtypes = [H]
    L.tkb_work.argtypes = [H, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64)]
    L.tkb_tc_time.argtypes = [H, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int)]
    L.tkb_trace_enable.argtypes = [H, ctypes.c_int]
    L.tkb_trace_sync.argtypes = [H]
    L.tkb_trace_save.argtypes = [H, ctypes.c_char_p]
    L.tkb_trace_blob.argtypes = [H, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_size_t)]
    L.tkb_trace_bytes.restype = ctypes.c_int64
    L.tkb_trace_bytes.argtypes = [H]
    L.tkb_trace_d2h_bytes.restype = ctypes.c_int64
    L.tkb_trace_d2h_bytes.argtypes = [H]
    L.tkb_fixed_point_multiplier_shift.argtypes = [ctypes.c_double, ctypes.POINTER(ctypes.c_int32),
                                                   ctypes.POINTER(ctypes.c_int32)]
    L.tkb_fixed_point_multiplier_shift.restype = None
    L.tkb_probe_int8_peak.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.c_void_p]
    _LIB = L
    return L


def _check(rc):
    if rc != 0:
        raise TachikomaError(load_library().tkb_last_error().decode())


def _is_torch(x):
    return type(x).__module__.startswith("torch")


class _Arg:
    """A DLTensor view of a numpy array (host) or torch tensor (host or CUDA); keeps the owner alive."""

    def __init__(self, x):
        if _is_torch(x):
            if not x.is_contiguous():
                raise TachikomaError("only compact tensors are supported (strides must be NULL)")
            self.owner = x
            dtype = str(x.dtype).replace("torch.", "")
            shape = tuple(x.shape)
            ptr = x.data_ptr()
            if x.is_cuda:
                dev = (kDLCUDA, x.device.index or 0)
            else:
                dev = (kDLCUDAHost if x.is_pinned() else kDLCPU, 0)
        else:
            x = np.asarray(x)
            if not x.flags["C_CONTIGUOUS"]:
                raise TachikomaError("only compact tensors are supported (strides must be NULL)")
            self.owner = x
            dtype, shape, ptr, dev = str(x.dtype), x.shape, x.ctypes.data, (kDLCPU, 0)
        code, bits = _DT[dtype]
        self.shape = (ctypes.c_int64 * max(len(shape), 1))(*shape)
        self.t = DLTensor(ctypes.c_void_p(ptr), DLDevice(*dev), len(shape), DLDataType(code, bits, 1), self.shape, None, 0)


def _tensor_array(xs):
    args = [_Arg(x) for x in xs]
    arr = (ctypes.POINTER(DLTensor) * max(len(args), 1))(*[ctypes.pointer(a.t) for a in args])
    return args, arr


class TachikomaJSONRuntime:
    """runtime.TachikomaJSONRuntimeCreate(symbol, graph_json, const_names)
    (src/runtime/contrib/tachikoma/tachikoma_json_runtime.cc:900-906)."""

    type_key = "tachikoma_json"

    def __init__(self, symbol, graph_json, const_names, _handle=None):
        self._lib = load_library()
        self._h = ctypes.c_void_p()
        if _handle is not None:
            self._h = _handle
        else:
            names = (ctypes.c_char_p * max(len(const_names), 1))(*[n.encode() for n in const_names])
            _check(self._lib.tkb_create(symbol.encode(), graph_json.encode(), names, len(const_names),
                                        ctypes.byref(self._h)))
        self._keep = []

    def __del__(self):
        try:
            if self._h:
                self._lib.tkb_free(self._h)
                self._h = None
        except Exception:
            pass

    # -- module functions (json_runtime.h:73-103) ------------------------------------------------
    def get_symbol(self):
        return self._lib.tkb_get_symbol(self._h).decode()

    def get_const_vars(self):
        names = ctypes.POINTER(ctypes.c_char_p)()
        n = ctypes.c_int()
        _check(self._lib.tkb_get_const_vars(self._h, ctypes.byref(names), ctypes.byref(n)))
        return [names[i].decode() for i in range(n.value)]

    def get_function(self, name):
        sym = self.get_symbol()
        if name == "get_symbol":
            return self.get_symbol
        if name == "get_const_vars":
            return self.get_const_vars
        if name == sym:
            return self.run
        if name == "__init_" + sym:
            return self.init
        return None  # PackedFunc(nullptr)

    __getitem__ = get_function

    @property
    def num_inputs(self):
        return self._lib.tkb_num_inputs(self._h)

    @property
    def num_outputs(self):
        return self._lib.tkb_num_outputs(self._h)

    def init(self, consts):
        """``__init_<symbol>``: bind the constants (const_names order) and build the plan."""
        args, arr = _tensor_array(consts)
        _check(self._lib.tkb_init(self._h, arr, len(args)))
        return 0

    def run(self, *tensors, stream=None, trace_mode=None):
        """``<symbol>``(inputs..., outputs...): outputs are caller-allocated.  ``trace_mode`` (0/1/2) overrides the
        handle's default trace mode for this call only (per-stream trace state, see tkb_run_traced)."""
        args, arr = _tensor_array(tensors)
        if trace_mode is None:
            _check(self._lib.tkb_run(self._h, arr, len(args), ctypes.c_void_p(stream or 0)))
        else:
            _check(self._lib.tkb_run_traced(self._h, arr, len(args), ctypes.c_void_p(stream or 0), int(trace_mode)))

    def packed_run(self, *tensors):
        """Call through the TVMBackendPackedCFunc-shaped entry (include/tvm/runtime/c_backend_api.h:49-51):
        args[0] = module handle (kTVMOpaqueHandle), args[1..] = DLTensor* (kTVMDLTensorHandle)."""
        args, _ = _tensor_array(tensors)

        class TVMValue(ctypes.Union):
            _fields_ = [("v_int64", ctypes.c_int64), ("v_float64", ctypes.c_double), ("v_handle", ctypes.c_void_p)]

        n = len(args) + 1
        vals = (TVMValue * n)()
        codes = (ctypes.c_int * n)()
        vals[0].v_handle = self._h
        codes[0] = 3
        for i, a in enumerate(args):
            vals[i + 1].v_handle = ctypes.cast(ctypes.pointer(a.t), ctypes.c_void_p)
            codes[i + 1] = 7
        ret, rcode = TVMValue(), ctypes.c_int(0)
        _check(self._lib.tkb_packed_run(vals, codes, n, ctypes.byref(ret), ctypes.byref(rcode), None))

    # -- persistence (json_runtime.h:105-135) ------------------------------------------------------
    def save_to_binary(self):
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        _check(self._lib.tkb_save_binary(self._h, ctypes.byref(p), ctypes.byref(n)))
        return ctypes.string_at(p.value, n.value)

    @classmethod
    def load_from_binary(cls, blob):
        L = load_library()
        h = ctypes.c_void_p()
        buf = ctypes.create_string_buffer(blob, len(blob))
        _check(L.tkb_load_binary(buf, len(blob), ctypes.byref(h)))
        return cls(None, None, None, _handle=h)

    # -- options / introspection ---------------------------------------------------------------------
    def set_option(self, key, value):
        _check(self._lib.tkb_set_option(self._h, key.encode(), str(value).encode()))

    def plan_summary(self):
        return self._lib.tkb_plan_summary(self._h).decode()

    def launch_count(self):
        return int(self._lib.tkb_launch_count(self._h))

    def work(self):
        macs, hbm = ctypes.c_int64(), ctypes.c_int64()
        self._lib.tkb_work(self._h, ctypes.byref(macs), ctypes.byref(hbm))
        return macs.value, hbm.value

    def tc_macs(self):
        """int8 MACs of the plan's tcgen05 launches (the ones tc_time() times)."""
        self._lib.tkb_tc_macs.restype = ctypes.c_int64
        self._lib.tkb_tc_macs.argtypes = [ctypes.c_void_p]
        return int(self._lib.tkb_tc_macs(self._h))

    def tc_time(self):
        """(milliseconds, launches) of the tcgen05 GEMM kernels in the last run (needs option time_tc=1)."""
        ms, n = ctypes.c_double(), ctypes.c_int()
        _check(self._lib.tkb_tc_time(self._h, ctypes.byref(ms), ctypes.byref(n)))
        return ms.value, n.value

    # -- trace -------------------------------------------------------------------------------------------
    def trace_enable(self, mode):
        _check(self._lib.tkb_trace_enable(self._h, int(mode)))

    def trace_sync(self, stream=None):
        if stream is None:
            _check(self._lib.tkb_trace_sync(self._h))
        else:
            _check(self._lib.tkb_trace_sync_on(self._h, ctypes.c_void_p(stream)))

    def trace_blob_view(self, stream=None):
        """(address, length) of the serialized trace in the runtime's pinned buffer: no copy.  Valid until the next
        traced run on the same stream."""
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        if stream is None:
            _check(self._lib.tkb_trace_blob(self._h, ctypes.byref(p), ctypes.byref(n)))
        else:
            _check(self._lib.tkb_trace_blob_on(self._h, ctypes.c_void_p(stream), ctypes.byref(p), ctypes.byref(n)))
        return p.value, n.value

    def trace_blob(self, stream=None):
        p, n = self.trace_blob_view(stream)
        return ctypes.string_at(p, n)

    def trace_dump_debug(self, directory):
        """Write output_tensors.params + _tvmdbg_graph_dump.json (debug_result.py:25, :114-149, :191-202)."""
        _check(self._lib.tkb_trace_dump_debug(self._h, str(directory).encode()))

    def trace_save(self, path):
        _check(self._lib.tkb_trace_save(self._h, path.encode()))

    def trace_bytes(self):
        return int(self._lib.tkb_trace_bytes(self._h))

    def trace_d2h_bytes(self):
        """Bytes copied device -> host per traced run (each distinct activation once)."""
        return int(self._lib.tkb_trace_d2h_bytes(self._h))


def fixed_point_multiplier_shift(m):
    s, e = ctypes.c_int32(), ctypes.c_int32()
    load_library().tkb_fixed_point_multiplier_shift(float(m), ctypes.byref(s), ctypes.byref(e))
    return s.value, e.value


def spin(microseconds, stream=None):
    """Enqueue a device busy-wait (measurement aid, see tkb_spin)."""
    L = load_library()
    L.tkb_spin.argtypes = [ctypes.c_double, ctypes.c_void_p]
    _check(L.tkb_spin(float(microseconds), ctypes.c_void_p(stream) if stream else None))


def probe_int8_peak(iters=2000):
    t = ctypes.c_double()
    _check(load_library().tkb_probe_int8_peak(int(iters), ctypes.byref(t), None))
    return t.value


class Conv2dDesc(ctypes.Structure):
    """tkb_conv2d_desc (include/tkb200.h)."""
    _fields_ = [(k, ctypes.c_int) for k in
                ("n", "c", "h", "w", "o", "kh", "kw", "stride_h", "stride_w", "dilation_h", "dilation_w", "pad_top",
                 "pad_left", "pad_bottom", "pad_right", "data_dtype_code", "weight_dtype_code")] + \
               [("input_zero_point", ctypes.c_int32), ("kernel_zero_point", ctypes.c_int32)]


#: static scratch of tkb_qnn_conv2d_noalloc (TKB_NOALLOC_SCRATCH_BYTES / TKB_NOALLOC_MAX_K in include/tkb200.h)
NOALLOC_SCRATCH_BYTES = 16 << 20
NOALLOC_MAX_K = 12288


def qnn_conv2d_noalloc(desc, data_ptr, weight_ptr, out_ptr, stream=None):
    """tkb_qnn_conv2d_noalloc: one qnn.conv2d on DEVICE pointers, no allocation inside the call (no counterpart in the
    reference; for callers that cannot cudaMalloc).  All calls share one static scratch array: serialize them."""
    L = load_library()
    L.tkb_qnn_conv2d_noalloc.argtypes = [ctypes.POINTER(Conv2dDesc), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.c_void_p]
    _check(L.tkb_qnn_conv2d_noalloc(ctypes.byref(desc), ctypes.c_void_p(data_ptr), ctypes.c_void_p(weight_ptr),
                                    ctypes.c_void_p(out_ptr), ctypes.c_void_p(stream) if stream else None))


def create_executor(func, params=None, mode="int", conv_impl="auto"):
    """Partition + codegen + create + init in one call; returns (module, run) where
    ``run(**inputs) -> numpy array`` uses host buffers (the reference's graph-executor usage)."""
    from . import tachikoma as T

    part = T.partition_for_tachikoma(func, params, mode=mode)
    symbol, graph_json, const_names, consts = T.serialize(part)
    mod = TachikomaJSONRuntime(symbol, graph_json, const_names)
    mod.set_option("conv_impl", conv_impl)
    mod.init(consts)
    out_shape, out_dtype = part.body.shape, part.body.dtype
    names = [p.name for p in part.params]
    dtypes = {p.name: p.dtype for p in part.params}
    shapes = {p.name: p.shape for p in part.params}

    def run(**inputs):
        ins = [np.asarray(np.asarray(inputs[n]).astype(dtypes[n]).reshape(shapes[n]), order="C") for n in names]
        out = np.empty(out_shape, dtype=out_dtype)
        mod.run(*ins, out)
        return out

    return mod, run
