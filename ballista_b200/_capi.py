"""ctypes binding of libballista_b200.so (include/ballista_b200.h).

This is the same call discipline a Rust `extern "C"` crate would use.  There is no CPU
fallback: if the CUDA library is missing, importing the operators fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libballista_b200.so")

# error codes (ballista_b200.h)
OK, ERR_INVALID_ARG, ERR_NOT_IMPLEMENTED, ERR_EXECUTION, ERR_ARROW, ERR_CUDA, ERR_NCCL, ERR_OOM = 0, -1, -2, -3, -4, -5, -6, -7

# b2_expr_kind / b2_binary_op / b2_dtype / misc enums
EXPR_COLUMN, EXPR_LITERAL, EXPR_BINARY, EXPR_NOT, EXPR_IS_NULL, EXPR_IS_NOT_NULL, EXPR_CAST, EXPR_NEGATIVE, EXPR_BETWEEN, EXPR_IN_LIST, EXPR_CASE, EXPR_SCALAR_FN = range(1, 13)
OP_AND, OP_OR, OP_EQ, OP_NOTEQ, OP_LTEQ, OP_LT, OP_GT, OP_GTEQ, OP_PLUS, OP_MINUS, OP_MULTIPLY, OP_DIVIDE, OP_LIKE = range(1, 14)
BOOL, INT32, INT64, UINT64, FLOAT64, DATE32, UTF8, UINT8 = range(1, 9)
# ScalarFunction (rust/core/proto/ballista.proto:80-114)
SCALAR_FUNCTIONS = ["sqrt", "sin", "cos", "tan", "asin", "acos", "atan", "exp", "log", "log2", "log10", "floor", "ceil",
                    "round", "trunc", "abs", "signum", "octetlength", "concat", "lower", "upper", "trim", "ltrim",
                    "rtrim", "totimestamp", "array", "nullif", "datetrunc", "md5", "sha224", "sha256", "sha384", "sha512"]
AGG_SUM, AGG_AVG, AGG_COUNT = 1, 2, 3
AGG_PARTIAL, AGG_FINAL = 0, 1
JOIN_INNER, JOIN_LEFT, JOIN_RIGHT = 0, 1, 2
PART_UNKNOWN, PART_ROUND_ROBIN, PART_HASH = 0, 1, 2


class BallistaError(RuntimeError):
    """Mirror of rust/core/src/error.rs:31-47 on the Python side of the ABI."""

    def __init__(self, code, msg):
        super().__init__(f"[{code}] {msg}")
        self.code = code


class NotImplementedOnGpu(BallistaError):
    pass


class ArrowSchema(C.Structure):
    pass


class ArrowArray(C.Structure):
    pass


ArrowSchema._fields_ = [
    ("format", C.c_char_p), ("name", C.c_char_p), ("metadata", C.c_char_p), ("flags", C.c_int64),
    ("n_children", C.c_int64), ("children", C.POINTER(C.POINTER(ArrowSchema))),
    ("dictionary", C.POINTER(ArrowSchema)), ("release", C.c_void_p), ("private_data", C.c_void_p),
]
ArrowArray._fields_ = [
    ("length", C.c_int64), ("null_count", C.c_int64), ("offset", C.c_int64), ("n_buffers", C.c_int64),
    ("n_children", C.c_int64), ("buffers", C.POINTER(C.c_void_p)), ("children", C.POINTER(C.POINTER(ArrowArray))),
    ("dictionary", C.POINTER(ArrowArray)), ("release", C.c_void_p), ("private_data", C.c_void_p),
]


class ExprNode(C.Structure):
    _fields_ = [("kind", C.c_int32), ("op", C.c_int32), ("dtype", C.c_int32), ("col", C.c_int32),
                ("nchildren", C.c_int32), ("reserved", C.c_int32), ("i", C.c_int64), ("f", C.c_double),
                ("s", C.c_char_p)]


class Expr(C.Structure):
    _fields_ = [("nodes", C.POINTER(ExprNode)), ("n_nodes", C.c_int32), ("name", C.c_char_p)]


class Agg(C.Structure):
    _fields_ = [("fn", C.c_int32), ("arg", Expr)]


_lib = None


def lib():
    """Load the CUDA library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(ballista_b200 has no CPU fallback)")
    try:
        import torch  # noqa: F401  (loads libnccl.so.2 / libcudart that the library links by soname)
    except Exception:
        pass
    L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    P = C.POINTER
    sigs = {
        "b2_last_error": (C.c_char_p, []),
        "b2_version": (C.c_char_p, []),
        "b2_ctx_create": (i32, [i32, P(vp)]),
        "b2_ctx_destroy": (None, [vp]),
        "b2_ctx_cuda_stream": (vp, [vp]),
        "b2_ctx_synchronize": (i32, [vp]),
        "b2_ctx_kernel_launches": (i64, [vp]),
        "b2_ctx_set_profiling": (i32, [vp, i32]),
        "b2_ctx_profile": (i32, [vp, P(C.c_double), P(i64)]),
        "b2_host_alloc": (i32, [vp, C.c_size_t, P(vp)]),
        "b2_host_free": (i32, [vp, vp]),
        "b2_memory_exec_create": (i32, [vp, P(ArrowSchema), i32, P(vp)]),
        "b2_memory_exec_add_batch": (i32, [vp, i32, P(ArrowArray)]),
        "b2_memory_exec_add_device_batch": (i32, [vp, i32, vp]),
        "b2_memory_exec_clear": (i32, [vp]),
        "b2_filter_create": (i32, [vp, vp, P(Expr), P(vp)]),
        "b2_coalesce_batches_create": (i32, [vp, vp, i64, P(vp)]),
        "b2_merge_create": (i32, [vp, vp, P(vp)]),
        "b2_projection_create": (i32, [vp, vp, P(Expr), i32, P(vp)]),
        "b2_csv_decode": (i32, [vp, P(ArrowSchema), P(C.c_int), i32, C.c_char, i32, C.c_char_p, C.c_size_t, P(vp)]),
        "b2_parquet_describe": (i32, [vp, C.c_size_t, P(i64), P(i32), P(i32), C.c_char_p, C.c_size_t]),
        "b2_parquet_schema": (i32, [vp, C.c_size_t, P(C.c_int), i32, P(ArrowSchema)]),
        "b2_parquet_decode": (i32, [vp, vp, C.c_size_t, P(C.c_int), i32, i32, i32, P(vp)]),
        "b2_plan_from_proto": (i32, [vp, C.c_char_p, C.c_size_t, P(vp), i32, P(vp)]),
        "b2_plan_proto_describe": (i32, [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, P(i32)]),
        "b2_jit_compile_check": (i32, [C.c_char_p, P(C.c_size_t)]),
        "b2_sort_create": (i32, [vp, vp, P(Expr), P(C.c_int), P(C.c_int), i32, P(vp)]),
        "b2_limit_create": (i32, [vp, vp, C.c_int64, i32, P(vp)]),
        "b2_hash_aggregate_create": (i32, [vp, vp, i32, P(Expr), i32, P(Agg), i32, P(ArrowSchema), P(vp)]),
        "b2_hash_join_create": (i32, [vp, vp, vp, P(C.c_char_p), P(C.c_char_p), i32, i32, P(vp)]),
        "b2_repartition_create": (i32, [vp, vp, i32, P(Expr), i32, i32, P(vp)]),
        "b2_op_schema": (i32, [vp, P(ArrowSchema)]),
        "b2_op_output_partitioning": (i32, [vp, P(i32), P(i32)]),
        "b2_op_n_children": (i32, [vp]),
        "b2_op_child": (vp, [vp, i32]),
        "b2_op_name": (C.c_char_p, [vp]),
        "b2_op_execute": (i32, [vp, i32, P(vp)]),
        "b2_op_destroy": (None, [vp]),
        "b2_op_last_path": (C.c_char_p, [vp]),
        "b2_stream_next": (i32, [vp, P(ArrowArray), P(i32)]),
        "b2_stream_next_device": (i32, [vp, P(vp), P(i32)]),
        "b2_stream_destroy": (None, [vp]),
        "b2_batch_from_arrow": (i32, [vp, P(ArrowSchema), P(ArrowArray), P(vp)]),
        "b2_batch_to_arrow": (i32, [vp, P(ArrowArray)]),
        "b2_batch_schema": (i32, [vp, P(ArrowSchema)]),
        "b2_batch_num_rows": (i64, [vp]),
        "b2_batch_num_bytes": (i64, [vp]),
        "b2_batch_destroy": (None, [vp]),
        "b2_tpch_generate": (i32, [vp, C.c_char_p, C.c_double, i64, i64, P(C.c_char_p), P(vp)]),
        "b2_tpch_num_rows": (i64, [C.c_char_p, C.c_double]),
        "b2_comm_unique_id": (i32, [vp]),
        "b2_comm_init": (i32, [vp, i32, i32, vp]),
        "b2_comm_destroy": (i32, [vp]),
        "b2_comm_allgather_batch": (i32, [vp, vp, P(vp)]),
        "b2_comm_bytes_sent": (i64, [vp]),
        "b2_hash_split_i32_static": (i32, [i32, vp, vp, i64, i32, vp, vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(L, name)  # AttributeError here = header/library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    L._b2_symbols = sorted(sigs)
    _lib = L
    return L


def check(code):
    if code == OK:
        return
    msg = lib().b2_last_error().decode("utf-8", "replace")
    if code == ERR_NOT_IMPLEMENTED:
        raise NotImplementedOnGpu(code, msg)
    raise BallistaError(code, msg)
