"""ballista_b200 — B200-native (sm_100a) executor-side operators for Ballista's hot path:
FilterExec, HashAggregateExec (Partial/Final), HashJoinExec, RepartitionExec (+ the in-box NVLink exchange).

All compute lives in libballista_b200.so (hand-written CUDA, C ABI in include/ballista_b200.h); this package
is the host-side mirror of the reference's operator interface.  There is no CPU fallback.
"""
from . import _capi
from ._capi import BallistaError, NotImplementedOnGpu
from .expr import Expr, AggExpr, col, lit, date, case, scalar_fn, sum, avg, count  # noqa: A004
from .plan import (Context, DeviceBatch, ExecutionPlan, RecordBatchStream, MemoryExec, FilterExec,
                   CoalesceBatchesExec, MergeExec, ProjectionExec, HashAggregateExec, HashJoinExec,
                   RepartitionExec, SortExec, GlobalLimitExec, LocalLimitExec, plan_from_proto, describe_plan_proto)
from .parquet import ParquetFile, ParquetExec
from . import tpch

__all__ = ["BallistaError", "NotImplementedOnGpu", "Expr", "AggExpr", "col", "lit", "date", "case", "scalar_fn", "sum", "avg", "count",
           "Context", "DeviceBatch", "ExecutionPlan", "RecordBatchStream", "MemoryExec", "FilterExec",
           "CoalesceBatchesExec", "MergeExec", "ProjectionExec", "HashAggregateExec", "HashJoinExec",
           "RepartitionExec", "SortExec", "GlobalLimitExec", "LocalLimitExec", "plan_from_proto",
           "describe_plan_proto", "ParquetFile", "ParquetExec", "tpch"]
