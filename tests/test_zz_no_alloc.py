"""GPU: b2_hash_split_i32_static — the allocation-free hash split (scratch in a file-scope device array of the library,
no counterpart in the reference; see include/ballista_b200.h) — against the existing entry point for the same work,
RepartitionExec(Partitioning::Hash) through b2_repartition_create / b2_op_execute.  Integer / index work: bit-exact.
This file sorts last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ROWS = 1 << 22


def _keys(n, seed):
    rng = np.random.default_rng(seed)
    k = rng.integers(-(1 << 31), (1 << 31) - 1, size=n, dtype=np.int64).astype(np.int32)
    k[: n // 4] = rng.integers(0, 1000, size=n // 4).astype(np.int32)  # duplicates: equal keys share a destination
    if n >= 4:
        k[-4:] = np.array([0, -1, np.iinfo(np.int32).min, np.iinfo(np.int32).max], dtype=np.int32)
    return k


def _split_static(lib, torch, keys_np, n_out):
    """one call of the new entry point on cuda:0 -> (perm, starts) as numpy; also asserts that it allocated nothing"""
    n = len(keys_np)
    dev = torch.device("cuda:0")
    keys = torch.from_numpy(keys_np).to(dev)
    perm = torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev)  # uint32 row ids, read back through a view
    starts = torch.full((n_out + 1,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    alloc0 = torch.cuda.memory_allocated(0)
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc = lib.b2_hash_split_i32_static(0, C.c_void_p(stream), C.c_void_p(keys.data_ptr()), n, n_out,
                                      C.c_void_p(perm.data_ptr()), C.c_void_p(starts.data_ptr()))
    assert rc == 0, lib.b2_last_error().decode()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert torch.cuda.memory_allocated(0) == alloc0
    return perm.cpu().numpy().view(np.uint32)[:n].astype(np.int64), starts.cpu().numpy(), (free0, free1)


def _split_existing(ctx, keys_np, n_out):
    """the existing entry point: RepartitionExec(Hash) over (k Int32, rowid Int64) -> row ids of every output partition"""
    import pyarrow as pa
    from ballista_b200 import RepartitionExec, col, tpch
    n = len(keys_np)
    # one input partition holding one batch: the split sees the rows in the order the new entry point is given
    b = pa.record_batch({"k": pa.array(keys_np), "rowid": pa.array(np.arange(n, dtype=np.int64))})
    rep = RepartitionExec(tpch.scan(ctx, [[b]]), "Hash", n_out, [col("k")])
    return [np.asarray(rep.collect(p)["rowid"].to_numpy(), dtype=np.int64) for p in range(n_out)]


def _check(ctx, lib, torch, n, n_out, seed):
    keys = _keys(n, seed)
    perm, starts, mem = _split_static(lib, torch, keys, n_out)
    want = _split_existing(ctx, keys, n_out)
    assert starts[0] == 0 and starts[n_out] == n and np.all(np.diff(starts) >= 0)
    for p in range(n_out):
        got = perm[starts[p]:starts[p + 1]]
        assert np.array_equal(got, want[p]), (n, n_out, p, len(got), len(want[p]))  # same rows, same order: bit-exact
    assert np.array_equal(np.sort(perm), np.arange(n))  # a permutation of the input rows
    return mem


def test_static_split_matches_repartition_exec_two_calls_in_a_row(ctx):
    import torch
    from ballista_b200 import _capi
    lib = _capi.lib()
    # first call: large, every tile and all 32 destinations used; second call right after it: smaller, fewer
    # destinations, a ragged last tile — must not see anything the first call left in the shared scratch array
    _check(ctx, lib, torch, 1_000_003, 32, seed=1)
    mem = _check(ctx, lib, torch, 70_001, 5, seed=2)
    assert mem[1] >= mem[0]  # the (warm) call took no device memory
    _check(ctx, lib, torch, 2047, 8, seed=3)
    _check(ctx, lib, torch, 1, 1, seed=4)


def test_static_split_at_the_limit_and_beyond(ctx):
    import torch
    from ballista_b200 import _capi
    lib = _capi.lib()
    _check(ctx, lib, torch, MAX_ROWS, 8, seed=5)
    # beyond the static scratch: refused with a clear message, nothing launched (outputs untouched)
    dev = torch.device("cuda:0")
    keys = torch.zeros(16, dtype=torch.int32, device=dev)
    perm = torch.zeros(16, dtype=torch.int32, device=dev)
    starts = torch.full((9,), -7, dtype=torch.int64, device=dev)
    rc = lib.b2_hash_split_i32_static(0, None, C.c_void_p(keys.data_ptr()), MAX_ROWS + 1, 8,
                                      C.c_void_p(perm.data_ptr()), C.c_void_p(starts.data_ptr()))
    assert rc == -1 and b"maximum of 4194304 rows" in lib.b2_last_error()
    rc = lib.b2_hash_split_i32_static(0, None, C.c_void_p(keys.data_ptr()), 16, 33,
                                      C.c_void_p(perm.data_ptr()), C.c_void_p(starts.data_ptr()))
    assert rc == -1 and b"1..32" in lib.b2_last_error()
    torch.cuda.synchronize()
    assert starts.cpu().tolist() == [-7] * 9
    # n_rows = 0: every start is 0
    rc = lib.b2_hash_split_i32_static(0, None, None, 0, 8, None, C.c_void_p(starts.data_ptr()))
    assert rc == 0, lib.b2_last_error().decode()
    torch.cuda.synchronize()
    assert starts.cpu().tolist() == [0] * 9
