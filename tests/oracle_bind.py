"""ctypes bindings of oracle/liboracle.so — the CPU restatement of the reference (TEST INFRASTRUCTURE).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import this."""
import ctypes
import json
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
_lib = None


def lib():
    global _lib
    if _lib is None:
        path = ROOT / "oracle" / "liboracle.so"
        if not path.exists():
            raise RuntimeError("oracle/liboracle.so is not built: run `make -C oracle`")
        o = ctypes.CDLL(str(path))
        vp, cp = ctypes.c_void_p, ctypes.c_char_p
        o.oracle_last_error.restype = cp
        o.oracle_free.argtypes = [vp]
        o.oracle_execute_query.restype = vp
        o.oracle_execute_query.argtypes = [vp, cp, ctypes.c_int, ctypes.c_int]
        o.oracle_eval_optree.argtypes = [cp, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
        o.oracle_filter_ids.argtypes = [vp, cp, ctypes.c_uint32, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
        o.oracle_filter_cardinality.argtypes = [vp, cp, ctypes.c_uint32, ctypes.POINTER(ctypes.c_uint64)]
        o.oracle_mutation_counts.argtypes = [vp, cp, cp, ctypes.c_int, ctypes.c_int, vp, ctypes.c_size_t]
        o.oracle_mutation_counts_subset.argtypes = [vp, cp, cp, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_uint32), ctypes.c_uint32, vp, ctypes.c_size_t]
        o.oracle_time_query_subset.restype = ctypes.c_double
        o.oracle_time_query_subset.argtypes = [vp, cp, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_uint32), ctypes.c_uint32, ctypes.POINTER(ctypes.c_uint64)]
        _lib = o
    return _lib


class OracleError(RuntimeError):
    def __init__(self, message):
        super().__init__(message)
        self.kind, _, self.message = message.partition(":")


def _err():
    return OracleError(lib().oracle_last_error().decode())


def _ids(ptr, n):
    try:
        return np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(ctypes.c_uint32)), shape=(max(n, 1),))[:n].copy()
    finally:
        lib().oracle_free(ptr)


def execute(db, query, faithful=True, threads=0):
    q = query if isinstance(query, str) else json.dumps(query)
    r = lib().oracle_execute_query(db.handle, q.encode(), int(faithful), threads)
    if not r:
        raise _err()
    try:
        return json.loads(ctypes.string_at(r))["queryResult"]
    finally:
        lib().oracle_free(r)


def execute_subset_text(db, query, partitions):
    """response text of `query` restricted to `partitions` (what one rank of a sharded database answers)"""
    q = query if isinstance(query, str) else json.dumps(query)
    arr = (ctypes.c_uint32 * max(len(partitions), 1))(*partitions)
    fn = lib().oracle_execute_query_subset
    fn.restype = ctypes.c_void_p
    fn.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_uint32), ctypes.c_uint32]
    r = fn(db.handle, q.encode(), arr, len(partitions))
    if not r:
        raise _err()
    try:
        return ctypes.string_at(r).decode()
    finally:
        lib().oracle_free(r)


def eval_optree(tree):
    ptr, n = ctypes.c_void_p(), ctypes.c_size_t()
    if lib().oracle_eval_optree(json.dumps(tree).encode(), ctypes.byref(ptr), ctypes.byref(n)) != 0:
        raise _err()
    return _ids(ptr, n.value)


def filter_ids(db, filter_expression, partition):
    ptr, n = ctypes.c_void_p(), ctypes.c_size_t()
    if lib().oracle_filter_ids(db.handle, json.dumps(filter_expression).encode(), partition, ctypes.byref(ptr), ctypes.byref(n)) != 0:
        raise _err()
    return _ids(ptr, n.value)


def filter_cardinality(db, filter_expression, partition):
    n = ctypes.c_uint64()
    if lib().oracle_filter_cardinality(db.handle, json.dumps(filter_expression).encode(), partition, ctypes.byref(n)) != 0:
        raise _err()
    return n.value


def mutation_counts(db, query, sequence_name, table_len, faithful=True, threads=0):
    out = np.zeros(table_len, dtype=np.uint32)
    rc = lib().oracle_mutation_counts(db.handle, json.dumps(query).encode(), sequence_name.encode(), int(faithful), threads, out.ctypes.data, table_len)
    if rc not in (0, 4):
        raise _err()
    return out, rc == 0


def time_query_subset(db, query, partitions, faithful=True, threads=0):
    """seconds the CPU restatement needs for `query` restricted to `partitions` (bounded bench sample)"""
    arr = (ctypes.c_uint32 * max(len(partitions), 1))(*partitions)
    rows = ctypes.c_uint64()
    seconds = lib().oracle_time_query_subset(db.handle, json.dumps(query).encode(), int(faithful), threads, arr, len(partitions), ctypes.byref(rows))
    if seconds < 0:
        raise _err()
    return seconds, rows.value


def time_query_sample(db, query, partitions, position_limit, threads=0):
    """seconds of the faithful CPU restatement for `query` on `partitions`, positions [0, position_limit) only"""
    arr = (ctypes.c_uint32 * max(len(partitions), 1))(*partitions)
    fn = lib().oracle_time_query_sample
    fn.restype = ctypes.c_double
    fn.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(ctypes.c_uint32), ctypes.c_uint32, ctypes.c_uint32]
    seconds = fn(db.handle, json.dumps(query).encode(), threads, arr, len(partitions), position_limit)
    if seconds < 0:
        raise _err()
    return seconds


def mutation_counts_subset(db, query, sequence_name, table_len, partitions, faithful=True, threads=0):
    """the table of `query` restricted to `partitions` (the others count as empty filters)"""
    arr = (ctypes.c_uint32 * max(len(partitions), 1))(*partitions)
    out = np.zeros(table_len, dtype=np.uint32)
    rc = lib().oracle_mutation_counts_subset(db.handle, json.dumps(query).encode(), sequence_name.encode(), int(faithful), threads, arr, len(partitions), out.ctypes.data, table_len)
    if rc not in (0, 4):
        raise _err()
    return out, rc == 0
