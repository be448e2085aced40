"""CPU: shapes outside the pre-built fused kernels compile with NVRTC for sm_100a from the embedded headers (no device
needed to compile), and shapes outside the fused vocabulary are refused with a reason."""
import ctypes as C

import pytest


def check(sig):
    from ballista_b200 import _capi as K
    n = C.c_size_t(0)
    rc = K.lib().b2_jit_compile_check(sig.encode(), C.byref(n))
    return rc, n.value, K.lib().b2_last_error().decode()


def test_unseen_shapes_compile_for_sm100a():
    for sig in ["p:lt(i0,LI0)|k:s0|a:f0,mul(f0,f1)",                          # Q1-like with `<`, one key
                "p:or(and(ge(f0,LF0),ne(s0,LS0)),eq(i0,LI0))|k:i1,s0|a:div(f1,f2),neg(f0),f64(i0)",
                "p:true|k:|a:add(f0,LF0)"]:
        rc, size, err = check(sig)
        assert rc == 0 and size > 50_000, (sig, err)


@pytest.mark.parametrize("sig", ["p:like(s0,LS0)|k:|a:f0",        # LIKE has no fused form
                                 "p:true|k:x3|a:f0",              # Int64 key column
                                 "p:lt(i0,LF0)|k:|a:f0",          # mixed classes (decode would have inserted a cast)
                                 "p:true|k:s0|a:i0",              # aggregate arguments are Float64
                                 "garbage"])
def test_shapes_outside_the_vocabulary_are_refused(sig):
    from ballista_b200 import _capi as K
    rc, size, err = check(sig)
    assert rc == K.ERR_NOT_IMPLEMENTED and size == 0 and err
