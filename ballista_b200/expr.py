"""Expression builder mirroring the reference's logical expression vocabulary (rust/core/proto/ballista.proto:14-45;
Python binding python/src/expression.rs): col(), lit(), arithmetic / comparison / boolean operators, between,
in_list, cast, and the aggregate constructors sum / avg / count (to_proto.rs:352-363).

An expression is lowered to the flat post-order IR of include/ballista_b200.h against an input schema.
"""
import ctypes as C

from . import _capi as K


class Expr:
    def __init__(self, kind, children=(), **kw):
        self.kind = kind
        self.children = list(children)
        self.kw = kw
        self.alias_name = None
        self._lowered = {}  # (column names, default name) -> (IR struct, keep-alive objects); expressions are immutable

    # ---- naming ----
    def alias(self, name):
        e = Expr(self.kind, self.children, **self.kw)
        e.alias_name = name
        return e

    # ---- operators ----
    def _bin(self, op, other, swap=False):
        other = other if isinstance(other, Expr) else lit(other)
        a, b = (other, self) if swap else (self, other)
        return Expr(K.EXPR_BINARY, [a, b], op=op)

    def __add__(self, o): return self._bin(K.OP_PLUS, o)
    def __radd__(self, o): return self._bin(K.OP_PLUS, o, True)
    def __sub__(self, o): return self._bin(K.OP_MINUS, o)
    def __rsub__(self, o): return self._bin(K.OP_MINUS, o, True)
    def __mul__(self, o): return self._bin(K.OP_MULTIPLY, o)
    def __rmul__(self, o): return self._bin(K.OP_MULTIPLY, o, True)
    def __truediv__(self, o): return self._bin(K.OP_DIVIDE, o)
    def __rtruediv__(self, o): return self._bin(K.OP_DIVIDE, o, True)
    def __lt__(self, o): return self._bin(K.OP_LT, o)
    def __le__(self, o): return self._bin(K.OP_LTEQ, o)
    def __gt__(self, o): return self._bin(K.OP_GT, o)
    def __ge__(self, o): return self._bin(K.OP_GTEQ, o)
    def __eq__(self, o): return self._bin(K.OP_EQ, o)  # noqa: builds an expression, like DataFusion's Expr::eq
    def __ne__(self, o): return self._bin(K.OP_NOTEQ, o)
    def __and__(self, o): return self._bin(K.OP_AND, o)
    def __or__(self, o): return self._bin(K.OP_OR, o)
    def __invert__(self): return Expr(K.EXPR_NOT, [self])
    def __neg__(self): return Expr(K.EXPR_NEGATIVE, [self])
    __hash__ = object.__hash__

    def between(self, lo, hi, negated=False):
        lo = lo if isinstance(lo, Expr) else lit(lo)
        hi = hi if isinstance(hi, Expr) else lit(hi)
        return Expr(K.EXPR_BETWEEN, [self, lo, hi], op=int(negated))

    def in_list(self, items, negated=False):
        items = [i if isinstance(i, Expr) else lit(i) for i in items]
        return Expr(K.EXPR_IN_LIST, [self] + items, op=int(negated))

    def cast(self, dtype):
        return Expr(K.EXPR_CAST, [self], dtype=dtype)

    def like(self, pattern):
        """SQL LIKE (% and _ wildcards) against a literal pattern (BinaryExprNode op "Like")."""
        return self._bin(K.OP_LIKE, pattern)

    def is_null(self): return Expr(K.EXPR_IS_NULL, [self])
    def is_not_null(self): return Expr(K.EXPR_IS_NOT_NULL, [self])

    # ---- lowering ----
    def _emit(self, names, out, keep):
        for c in self.children:
            c._emit(names, out, keep)
        n = K.ExprNode()
        n.kind = self.kind
        n.op = self.kw.get("op", 0)
        n.dtype = self.kw.get("dtype", 0)
        n.col = -1
        n.nchildren = len(self.children)
        if self.kind == K.EXPR_COLUMN:
            name = self.kw["name"]
            if name not in names:
                raise K.BallistaError(K.ERR_INVALID_ARG, f"No field named '{name}'. Valid fields are {names}")
            n.col = names.index(name)
        elif self.kind == K.EXPR_LITERAL:
            v = self.kw["value"]
            if n.dtype == K.FLOAT64:
                n.f = float(v)
            elif n.dtype == K.UTF8:
                b = v.encode("utf-8")
                keep.append(b)
                n.s = b
            else:
                n.i = int(v)
        out.append(n)

    def lower(self, names, keep, default_name=None):
        """-> _capi.Expr against the column `names`; `keep` receives objects that must outlive the call.
        The lowered IR is cached on the expression (the library copies it at operator construction)."""
        key = (tuple(names), default_name)
        hit = self._lowered.get(key)
        if hit is None:
            mine, nodes = [], []
            self._emit(list(names), nodes, mine)
            arr = (K.ExprNode * len(nodes))(*nodes)
            mine.append(arr)
            nm = self.alias_name or default_name or self.display()
            b = nm.encode("utf-8")
            mine.append(b)
            hit = (C.cast(arr, C.POINTER(K.ExprNode)), len(nodes), b, mine)
            self._lowered[key] = hit
        e = K.Expr()
        e.nodes, e.n_nodes, e.name = hit[0], hit[1], hit[2]
        keep.append(hit[3])
        return e

    def display(self):
        if self.kind == K.EXPR_COLUMN:
            return self.kw["name"]
        if self.kind == K.EXPR_LITERAL:
            return repr(self.kw["value"])
        return f"expr{self.kind}({','.join(c.display() for c in self.children)})"


def col(name):
    return Expr(K.EXPR_COLUMN, name=name)


def lit(value, dtype=None):
    if dtype is None:
        if isinstance(value, bool):
            dtype = K.BOOL
        elif isinstance(value, int):
            dtype = K.INT64  # SQL integer literals are Int64 (ScalarValue::Int64)
        elif isinstance(value, float):
            dtype = K.FLOAT64
        elif isinstance(value, str):
            dtype = K.UTF8
        else:
            raise TypeError(f"unsupported literal {value!r}")
    return Expr(K.EXPR_LITERAL, value=value, dtype=dtype)


def case(when_then, otherwise=None, base=None):
    """CASE [base] WHEN w THEN t ... [ELSE otherwise] END (CaseNode, ballista.proto:141-150).
    `when_then` = [(when, then), ...]."""
    wrap = lambda v: v if isinstance(v, Expr) else lit(v)
    children = [wrap(base)] if base is not None else []
    for w, t in when_then:
        children += [wrap(w), wrap(t)]
    if otherwise is not None:
        children.append(wrap(otherwise))
    return Expr(K.EXPR_CASE, children, op=(1 if base is not None else 0) | (2 if otherwise is not None else 0))


def scalar_fn(name, *args):
    """ScalarFunctionNode (ballista.proto:80-119): sqrt, sin, ..., floor, ceil, round, trunc, abs, signum,
    octetlength, lower, upper, trim, ltrim, rtrim, nullif, md5, concat."""
    return Expr(K.EXPR_SCALAR_FN, [a if isinstance(a, Expr) else lit(a) for a in args],
                op=K.SCALAR_FUNCTIONS.index(name.lower()))


def date(s):
    """SQL `date 'YYYY-MM-DD'` = CAST(Utf8 AS Date32)."""
    return lit(s).cast(K.DATE32)


class AggExpr:
    """AggregateExprNode: fn in {SUM, AVG, COUNT}, one argument, non-distinct."""

    def __init__(self, fn, arg, name=None):
        self.fn, self.arg, self.name = fn, arg, name

    def alias(self, name):
        return AggExpr(self.fn, self.arg, name)


def sum(e): return AggExpr(K.AGG_SUM, e)  # noqa: A001 (mirrors datafusion::logical_plan::sum)
def avg(e): return AggExpr(K.AGG_AVG, e)
def count(e=None): return AggExpr(K.AGG_COUNT, e if e is not None else lit(1, K.UINT8))  # COUNT(*) = COUNT(UInt8(1))
