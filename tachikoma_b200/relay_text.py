"""Relay TEXT format (the reference's `astext()` / `tvm.parser.fromtext`) for the operators of the integer path
(first half).

    text, constants = astext(func)          # mini-IR Function -> Relay text + the meta[relay.Constant] arrays
    func = fromtext(text, constants)        # Relay text (as printed by TVM's RelayTextPrinter) -> mini-IR Function

Printing follows src/printer/relay_text_printer.cc: A-normal form with `%i = op(args, key=value, ...);`, rank-0 int32 /
float32 / ... constants inline with the suffixes of src/support/scalars.cc:95-124 (`5`, `5i64`, `0.5f`, `True`), every
other constant as `meta[relay.Constant][i]` (the arrays travel separately, like TVM's metadata section or a params
file), `/* ty=... */` annotations are comments.  One deliberate difference: float32 scalars are printed with nine
significant digits (TVM prints six, scalars.cc:111-113, which does not round-trip a quantization scale); the parser
accepts both.  Only the constructs the integer path needs are parsed: one function (`def @main(...)` or `fn (...)`),
calls of the supported operators, tuples as arguments of qnn.concatenate.
"""
from __future__ import annotations

import re

import numpy as np

from . import relay as R

_SUFFIX = {"int16": "i16", "int32": "", "int64": "i64", "float16": "f16", "float32": "f", "float64": "f64"}


# ----------------------------------------------------------------------------------------------------------------
# printer
# ----------------------------------------------------------------------------------------------------------------
def _scalar_text(a):
    dt = str(a.dtype)
    v = a.reshape(()).item()
    if dt == "bool":
        return "True" if v else "False"
    if dt.startswith("float"):
        s = "%.9g" % v if dt == "float32" else repr(float(v))
        return s + _SUFFIX[dt]
    return "%d%s" % (v, _SUFFIX[dt])


def _attr_text(v):
    if v is None:
        return "None"
    if isinstance(v, bool):
        return "True" if v else "False"
    if isinstance(v, (int, np.integer)):
        return str(int(v))
    if isinstance(v, (float, np.floating)):
        return "%.17g" % float(v) + "f"
    if isinstance(v, str):
        return '"%s"' % v
    if isinstance(v, (list, tuple)):
        return "[" + ", ".join(_attr_text(x) for x in v) + "]"
    raise TypeError("attribute of type %s" % type(v))


def _type_text(shape, dtype):
    dims = ", ".join(str(int(d)) for d in shape)
    if len(shape) == 1:
        dims += ","
    return "Tensor[(%s), %s]" % (dims, dtype) if len(shape) else dtype


def astext(func, annotate_types=True):
    """-> (text, constants): constants[i] is the array behind meta[relay.Constant][i]."""
    consts, const_id, names, lines = [], {}, {}, []

    def ref(e, scalar_ok=False):
        if isinstance(e, R.Var):
            return "%" + e.name
        if isinstance(e, R.Constant):
            a = np.asarray(e.data)
            # the mini IR stores scalars as one-element vectors (np.ascontiguousarray): in a quantization-parameter or
            # pad-value position they are the reference's rank-0 constants, which print inline
            if (a.ndim == 0 or (scalar_ok and a.shape == (1,))) and (str(a.dtype) in _SUFFIX or a.dtype == np.bool_):
                return _scalar_text(a)
            if id(e) not in const_id:
                const_id[id(e)] = len(consts)
                consts.append(a)
            t = " /* ty=%s */" % _type_text(a.shape, a.dtype) if annotate_types else ""
            return "meta[relay.Constant][%d]%s" % (const_id[id(e)], t)
        return names[id(e)]

    order = [e for e in R.post_order(func.body) if isinstance(e, R.Call)]
    for e in order:
        a = e.args
        if e.op == "qnn.concatenate":
            n = e.attrs["num_inputs"]
            tup = lambda xs, ok: "(" + ", ".join(ref(x, ok) for x in xs) + ("," if len(xs) == 1 else "") + ")"
            arg_text = [tup(a[:n], False), tup(a[n:2 * n], True), tup(a[2 * n:3 * n], True), ref(a[3 * n], True), ref(a[3 * n + 1], True)]
            attrs = {"axis": e.attrs["axis"]}
        else:
            first_param = 2 if e.op.startswith("qnn.") and e.op not in ("qnn.quantize", "qnn.dequantize", "qnn.requantize",
                                                                        "qnn.leaky_relu") + R.QNN_UNARY_OPS else 1
            scalar_pos = e.op.startswith("qnn.") or e.op == "nn.pad"
            arg_text = [ref(x, scalar_pos and i >= first_param) for i, x in enumerate(a)]
            attrs = dict(e.attrs)
        body = "%s(%s)" % (e.op, ", ".join(arg_text + ["%s=%s" % (k, _attr_text(v)) for k, v in attrs.items()]))
        ty = " /* ty=%s */" % _type_text(e.shape, e.dtype) if annotate_types else ""
        if e is func.body:
            lines.append("  %s%s" % (body, ty))
        else:
            names[id(e)] = "%%%d" % len(names)
            lines.append("  %s = %s%s;" % (names[id(e)], body, ty))
    if not order:
        lines.append("  " + ref(func.body))
    params = ", ".join("%%%s: %s" % (p.name, _type_text(p.shape, p.dtype)) for p in func.params)
    head = "def @main(%s) -> %s {" % (params, _type_text(func.body.shape, func.body.dtype))
    return '#[version = "0.0.5"]\n' + head + "\n" + "\n".join(lines) + "\n}\n", consts


# ----------------------------------------------------------------------------------------------------------------
# parser
# ----------------------------------------------------------------------------------------------------------------
_TOKEN = re.compile(r"""
    (?P<ws>\s+)
  | (?P<meta>meta\[relay\.Constant\]\[\d+\])
  | (?P<str>"[^"]*")
  | (?P<num>[-+]?(?:(?:\d+\.\d*|\.\d+|\d+)(?:[eE][-+]?\d+)?|inf|nan)(?:f64|f16|f|i64|i16|i8|u8)?(?![A-Za-z0-9_.]))
  | (?P<local>%[A-Za-z0-9_.]+)
  | (?P<glob>@[A-Za-z0-9_.]+)
  | (?P<name>[A-Za-z_][A-Za-z0-9_.]*)
  | (?P<arrow>->)
  | (?P<sym>[()\[\]{},;:=])
""", re.X)


def _lex(text):
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)  # type / span annotations are comments
    text = re.sub(r"^\s*#\[.*?\]\s*$", " ", text, flags=re.M)  # #[version = "0.0.5"], #[metadata] is not supported
    out, pos = [], 0
    while pos < len(text):
        m = _TOKEN.match(text, pos)
        if not m:
            raise ValueError("Relay text: cannot tokenize at %r" % text[pos:pos + 30])
        pos = m.end()
        if m.lastgroup != "ws":
            out.append((m.lastgroup, m.group()))
    return out


class _Parser:
    def __init__(self, toks, constants):
        self.t, self.i, self.consts, self.env = toks, 0, constants, {}

    def peek(self, k=0):
        return self.t[self.i + k] if self.i + k < len(self.t) else ("eof", "")

    def next(self):
        tok = self.peek()
        self.i += 1
        return tok

    def expect(self, val):
        kind, v = self.next()
        if v != val:
            raise ValueError("Relay text: expected %r, got %r" % (val, v))

    def accept(self, val):
        if self.peek()[1] == val:
            self.i += 1
            return True
        return False

    # ---- types ----
    def parse_type(self):
        kind, v = self.next()
        if v == "Tensor":
            self.expect("[")
            self.expect("(")
            dims = []
            while not self.accept(")"):
                dims.append(int(self.next()[1]))
                self.accept(",")
            self.expect(",")
            dtype = self.next()[1]
            self.expect("]")
            return tuple(dims), dtype
        return (), v  # scalar type such as int32

    # ---- literals ----
    @staticmethod
    def number(tok):
        m = re.match(r"^([-+]?(?:(?:\d+\.\d*|\.\d+|\d+)(?:[eE][-+]?\d+)?|inf|nan))(f64|f16|f|i64|i16|i8|u8)?$", tok)
        body, suf = m.group(1), m.group(2)
        if suf in ("f", "f16", "f64") or (suf is None and re.search(r"[.eE]|inf|nan", body)):
            dt = {"f": "float32", "f16": "float16", "f64": "float64", None: "float32"}[suf]
            return np.asarray(float(body), dtype=dt)
        dt = {None: "int32", "i64": "int64", "i16": "int16", "i8": "int8", "u8": "uint8"}[suf]
        return np.asarray(int(body), dtype=dt)

    def attr_value(self):
        kind, v = self.next()
        if v == "[":
            out = []
            while not self.accept("]"):
                out.append(self.attr_value())
                self.accept(",")
            return out
        if kind == "str":
            return v[1:-1]
        if kind == "num":  # attribute floats are doubles in the reference (the `f` suffix is only notation here)
            m = re.match(r"^(.*?)(f64|f16|f|i64|i16|i8|u8)?$", v)
            body, suf = m.group(1), m.group(2)
            return float(body) if (suf in ("f", "f16", "f64") or re.search(r"[.eE]|inf|nan", body)) else int(body)
        if v in ("True", "False"):
            return v == "True"
        if v == "None":
            return None
        raise ValueError("Relay text: attribute value %r" % v)

    # ---- expressions ----
    def operand(self):
        kind, v = self.peek()
        if v == "(":  # tuple
            self.next()
            items = []
            while not self.accept(")"):
                items.append(self.operand())
                self.accept(",")
            return items
        self.next()
        if kind == "local":
            if v not in self.env:
                raise ValueError("Relay text: %s used before its definition" % v)
            return self.env[v]
        if kind == "meta":
            idx = int(re.findall(r"\[(\d+)\]$", v)[0])
            if idx >= len(self.consts):
                raise ValueError("Relay text: meta[relay.Constant][%d] but only %d constants given" % (idx, len(self.consts)))
            key = ("meta", idx)
            if key not in self.env:
                self.env[key] = R.Constant(np.asarray(self.consts[idx]))
            return self.env[key]
        if kind == "num":
            return R.Constant(self.number(v))
        if v in ("True", "False"):
            return R.Constant(np.asarray(v == "True"))
        raise ValueError("Relay text: operand %r" % v)

    def call(self):
        kind, op = self.next()
        if kind != "name":
            raise ValueError("Relay text: operator name expected, got %r" % op)
        self.expect("(")
        args, attrs = [], {}
        while not self.accept(")"):
            if self.peek()[0] == "name" and self.peek(1)[1] == "=":
                k = self.next()[1]
                self.next()
                attrs[k] = self.attr_value()
            else:
                args.append(self.operand())
            self.accept(",")
        return _build(op, args, attrs)

    def expr(self):
        kind, v = self.peek()
        if kind == "name" and self.peek(1)[1] == "(":
            return self.call()
        return self.operand()

    def function(self):
        kind, v = self.next()
        if v == "def":
            self.next()  # @main
        elif v != "fn":
            raise ValueError("Relay text: expected `def @main` or `fn`, got %r" % v)
        self.expect("(")
        params = []
        while not self.accept(")"):
            name = self.next()[1]
            self.expect(":")
            shape, dtype = self.parse_type()
            var = R.var(name[1:], shape, dtype)
            self.env[name] = var
            params.append(var)
            self.accept(",")
        if self.accept("->"):
            self.parse_type()
        self.expect("{")
        body = None
        while True:
            if self.peek()[0] == "local" and self.peek(1)[1] == "=":
                name = self.next()[1]
                self.next()
                self.env[name] = self.expr()
                self.expect(";")
                continue
            body = self.expr()
            self.accept(";")
            break
        self.expect("}")
        return R.Function(params, body)


def _pick(attrs, *names):
    return {k: attrs[k] for k in names if k in attrs and attrs[k] is not None}


def _build(op, a, at):
    """Operator name + parsed operands + attributes -> mini-IR call (shapes and dtypes come from the builders)."""
    if op == "qnn.conv2d":
        w = a[1]
        lay = at.get("kernel_layout", "OIHW")
        ks = at.get("kernel_size") or [w.shape[lay.index("H")], w.shape[lay.index("W")]]
        ch = at.get("channels") or w.shape[lay.index("O")]
        return R.qnn_conv2d(a[0], w, a[2], a[3], a[4], a[5], ks, ch, **_pick(at, "strides", "padding", "dilation", "groups",
                                                                            "data_layout", "kernel_layout", "out_layout"),
                            out_dtype=at.get("out_dtype") or "int32")
    if op == "qnn.dense":
        return R.qnn_dense(a[0], a[1], a[2], a[3], a[4], a[5], at.get("units") or a[1].shape[0], out_dtype=at.get("out_dtype") or "int32")
    if op == "qnn.batch_matmul":
        return R.qnn_batch_matmul(*a[:6], out_dtype=at.get("out_dtype") or "int32")
    if op == "qnn.requantize":
        return R.qnn_requantize(*a[:5], **_pick(at, "axis", "rounding", "compute_dtype", "out_dtype"))
    if op in ("qnn.add", "qnn.mul", "qnn.subtract"):
        fn = {"qnn.add": R.qnn_add, "qnn.mul": R.qnn_mul, "qnn.subtract": R.qnn_subtract}[op]
        return fn(*a[:8], **_pick(at, "lhs_axis", "rhs_axis"))
    if op == "qnn.concatenate":
        return R.qnn_concatenate(a[0], a[1], a[2], a[3], a[4], at["axis"])
    if op == "qnn.quantize":
        return R.qnn_quantize(*a[:3], **_pick(at, "axis", "out_dtype"))
    if op == "qnn.dequantize":
        return R.qnn_dequantize(*a[:3], **_pick(at, "axis"))
    if op == "qnn.leaky_relu":
        return R.qnn_leaky_relu(a[0], at["alpha"], *a[1:5])
    if op in R.QNN_UNARY_OPS:
        return R.qnn_unary(op, *a[:5])
    if op == "nn.bias_add":
        return R.bias_add(a[0], a[1], **_pick(at, "axis"))
    if op == "add":
        return R.add(a[0], a[1])
    if op == "clip":
        return R.clip(a[0], at["a_min"], at["a_max"])
    if op == "cast":
        return R.cast(a[0], at["dtype"])
    if op == "reinterpret":
        return R.reinterpret(a[0], at["dtype"])
    if op == "take":
        return R.take(a[0], a[1], at.get("axis", 0) or 0, at.get("mode", "fast"))
    if op in ("nn.max_pool2d", "nn.avg_pool2d"):
        if at.get("ceil_mode"):
            raise ValueError("Unsupported op: %s with ceil_mode" % op)
        kw = _pick(at, "pool_size", "strides", "padding", "layout")
        if op == "nn.avg_pool2d":
            return R.avg_pool2d(a[0], count_include_pad=bool(at.get("count_include_pad", False)), **kw)
        return R.max_pool2d(a[0], **kw)
    if op == "nn.global_avg_pool2d":
        return R.global_avg_pool2d(a[0], **_pick(at, "layout"))
    if op == "nn.pad":
        pw = at["pad_width"]
        val = float(np.asarray(a[1].data).reshape(-1)[0]) if len(a) > 1 else float(at.get("pad_value", 0.0))
        return R.pad(a[0], pw, val, at.get("pad_mode", "constant"))
    if op == "reshape":
        return R.reshape(a[0], at["newshape"])
    if op == "nn.batch_flatten":
        return R.batch_flatten(a[0])
    if op == "transpose":
        return R.transpose(a[0], at["axes"])
    # float operators of the tachikoma float patterns
    conv_kw = ("strides", "padding", "dilation", "groups", "channels", "kernel_size", "data_layout", "kernel_layout", "out_layout")
    if op in ("nn.conv1d", "nn.conv2d", "nn.conv3d"):
        fn = {"nn.conv1d": R.conv1d, "nn.conv2d": R.conv2d, "nn.conv3d": R.conv3d}[op]
        return fn(a[0], a[1], **_pick(at, *conv_kw), out_dtype=at.get("out_dtype") or "")
    if op in ("nn.conv2d_transpose", "nn.conv3d_transpose"):
        fn = R.conv2d_transpose if op == "nn.conv2d_transpose" else R.conv3d_transpose
        return fn(a[0], a[1], **_pick(at, *conv_kw, "output_padding"), out_dtype=at.get("out_dtype") or "")
    if op == "nn.dense":
        return R.dense(a[0], a[1], at.get("units"), out_dtype=at.get("out_dtype") or "")
    if op == "nn.batch_matmul":
        return R.batch_matmul(a[0], a[1], out_dtype=at.get("out_dtype") or "", transpose_a=bool(at.get("transpose_a", False)),
                              transpose_b=bool(at.get("transpose_b", True)))
    if op == "nn.batch_norm":
        return R.batch_norm(*a[:5], **_pick(at, "axis", "epsilon", "center", "scale"))
    if op == "nn.layer_norm":
        return R.layer_norm(*a[:3], **_pick(at, "axis", "epsilon", "center", "scale"))
    if op == "nn.softmax":
        return R.softmax(a[0], **_pick(at, "axis"))
    if op == "nn.leaky_relu":
        return R.leaky_relu(a[0], at.get("alpha", 0.01))
    unary = {"nn.relu": R.relu, "tanh": R.tanh, "sigmoid": R.sigmoid, "abs": R.abs_, "exp": R.exp, "log": R.log, "sqrt": R.sqrt,
             "round": R.round_, "erf": R.erf}
    if op in unary:
        return unary[op](a[0])
    if op in ("multiply", "divide"):
        return (R.multiply if op == "multiply" else R.divide)(a[0], a[1])
    if op in ("nn.max_pool3d", "nn.avg_pool3d"):
        if at.get("ceil_mode"):
            raise ValueError("Unsupported op: %s with ceil_mode" % op)
        kw = _pick(at, "pool_size", "strides", "dilation", "padding", "layout")
        if op == "nn.avg_pool3d":
            return R.avg_pool3d(a[0], count_include_pad=bool(at.get("count_include_pad", False)), **kw)
        return R.max_pool3d(a[0], **kw)
    raise ValueError("Unsupported op: %s" % op)


def fromtext(text, constants=()):
    """Relay text -> mini-IR Function.  `constants`: the arrays behind meta[relay.Constant][i], in index order."""
    p = _Parser(_lex(text), list(constants))
    f = p.function()
    if p.peek()[0] != "eof":
        raise ValueError("Relay text: trailing input %r (only one function is supported)" % (p.peek()[1],))
    return f
