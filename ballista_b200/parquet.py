"""ParquetExec leaf: Parquet files decoded on the GPU (csrc/parquet.cu, csrc/parquet_meta.cpp).

The reference builds this leaf as `ParquetExec::try_from_files(filenames, projection, None, batch_size, num_partitions)`
(rust/core/src/serde/physical_plan/from_proto.rs:111-121; message `ParquetScanExecNode { filename, projection,
num_partitions, batch_size }`, rust/core/proto/ballista.proto:325-330).  DataFusion@46161d2 splits the file list into
`num_partitions` contiguous chunks of ceil(n_files / num_partitions) files, one output partition per chunk; each file's
row groups are read in order.  `ParquetExec` here keeps exactly that partitioning; one batch per file (per ~3 GB of file)
instead of `batch_size`-row batches (the operators above take whole batches; CoalesceBatchesExec is a no-op on them).
"""
import ctypes as C
import mmap

import numpy as np
import pyarrow as pa

from . import _capi as K
from .plan import DeviceBatch, MemoryExec, _addr


class ParquetFile:
    """One Parquet file mapped into host memory; metadata through b2_parquet_describe / b2_parquet_schema (no GPU)."""

    def __init__(self, source, pinned_by=None):
        """source: a path or the file's bytes.  pinned_by=ctx reads the file into page-locked memory of that context
        (b2_host_alloc), which lets the column chunks cross PCIe at the link rate instead of the pageable-copy rate."""
        self._map = self._pinned = None
        self._ctx = pinned_by
        if pinned_by is not None:
            if isinstance(source, (bytes, bytearray, memoryview)):
                n = len(source)
            else:
                import os
                n = os.path.getsize(source)
            p = C.c_void_p()
            K.check(K.lib().b2_host_alloc(pinned_by._h, max(n, 1), C.byref(p)))
            self._pinned = p
            self._buf = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(max(n, 1),))[:n]
            if isinstance(source, (bytes, bytearray, memoryview)):
                self._buf[:] = np.frombuffer(source, dtype=np.uint8)
            else:
                with open(source, "rb") as f:
                    got = f.readinto(memoryview(self._buf))
                    assert got == n
        elif isinstance(source, (bytes, bytearray, memoryview)):
            self._buf = np.frombuffer(bytes(source), dtype=np.uint8)
        else:
            with open(source, "rb") as f:
                self._map = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ)
            self._buf = np.frombuffer(self._map, dtype=np.uint8)
        self._len = len(self._buf)
        rows, groups, cols = C.c_int64(0), C.c_int(0), C.c_int(0)
        K.check(K.lib().b2_parquet_describe(self._addr(), self._len, C.byref(rows), C.byref(groups), C.byref(cols), None, 0))
        self.num_rows, self.num_row_groups, self.num_columns = rows.value, groups.value, cols.value

    def __del__(self):
        try:
            if self._pinned is not None and self._ctx is not None and self._ctx._h:
                self._buf = None
                K.lib().b2_host_free(self._ctx._h, self._pinned)
                self._pinned = None
        except Exception:
            pass

    def _addr(self):
        return C.c_void_p(self._buf.ctypes.data)

    def describe(self):
        cap = 1 << 20
        out = C.create_string_buffer(cap)
        K.check(K.lib().b2_parquet_describe(self._addr(), self._len, None, None, None, out, cap))
        return out.value.decode("utf-8", "replace")

    def schema(self, projection=None):
        sch = K.ArrowSchema()
        arr, n = self._projection(projection)
        K.check(K.lib().b2_parquet_schema(self._addr(), self._len, arr, n, C.byref(sch)))
        return pa.Schema._import_from_c(_addr(sch))

    def _projection(self, projection):
        if projection is None:
            return None, 0
        proj = [int(p) for p in projection]
        return (C.c_int * max(len(proj), 1))(*proj), len(proj)

    def column_index(self, name):
        return self.schema().names.index(name)

    def read(self, ctx, projection=None, row_groups=None):
        """-> DeviceBatch of row groups [begin, end) (default: all) of the projected columns (indices or names)."""
        if projection is not None:
            names = None
            proj = []
            for p in projection:
                if isinstance(p, str):
                    names = names or self.schema().names
                    p = names.index(p)
                proj.append(p)
            projection = proj
        begin, end = (0, -1) if row_groups is None else row_groups
        arr, n = self._projection(projection)
        h = C.c_void_p()
        K.check(K.lib().b2_parquet_decode(ctx._h, self._addr(), self._len, arr, n, int(begin), int(end), C.byref(h)))
        return DeviceBatch(ctx, h)


def _read_batches(self, ctx, projection=None, max_bytes=3 << 30):
    """The whole file as few DeviceBatches as the 4 GiB-per-call limit allows (normally one): the pages of all row
    groups decode side by side, so the longest page chain is paid once per call, not once per row group."""
    if self.num_row_groups == 0:
        return [self.read(ctx, projection)]
    per_group = max(self._len // max(self.num_row_groups, 1), 1)
    step = max(int(max_bytes // per_group), 1)
    return [self.read(ctx, projection, (g, min(g + step, self.num_row_groups)))
            for g in range(0, self.num_row_groups, step)]


ParquetFile.read_batches = _read_batches


def split_files(filenames, n):
    """DataFusion's split_files: contiguous chunks of ceil(len / n) files."""
    filenames = list(filenames)
    if not filenames:
        return []
    size = len(filenames) // n + (1 if len(filenames) % n else 0)
    return [filenames[i:i + size] for i in range(0, len(filenames), size)]


class ParquetExec(MemoryExec):
    """ParquetExec::try_from_files(filenames, projection, None, batch_size, num_partitions): a leaf whose partition p
    holds the decoded row groups of the p-th chunk of files.  Decoding happens here, once; the plans above re-execute
    from HBM."""

    def __init__(self, ctx, filenames, projection=None, batch_size=32768, num_partitions=1):
        if not filenames:
            raise K.BallistaError("No files found")  # ParquetExec::try_from_files on an empty list
        chunks = split_files(filenames, max(int(num_partitions), 1))
        first = ParquetFile(filenames[0])
        schema = first.schema(projection)
        parts = []
        for chunk in chunks:
            batches = []
            for name in chunk:
                f = first if name == filenames[0] else ParquetFile(name)
                if f.schema(projection) != schema:
                    raise K.BallistaError(f"Parquet file {name} has a different schema")
                batches.extend(f.read_batches(ctx, projection))
            parts.append(batches)
        self.filenames, self.projection, self.batch_size = list(filenames), projection, batch_size
        super().__init__(ctx, schema, parts)
