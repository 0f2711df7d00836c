"""ctypes bindings of the product libraries (libsilo_host.so -> libsilo_b200.so).

Python is only the harness here (tests, bench.py, smoke): the host logic is C++ (host/*.hpp) and the
bitmap work is CUDA (csrc/). Nothing in this module computes a query result; a missing library or a
missing GPU raises.
"""
import ctypes
import json
import os
import subprocess
from pathlib import Path

import numpy as np

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
# kernel experiments (tools/k1_experiment.py) load a variant build of both libraries from another directory
LIBDIR = Path(os.environ["SILO_B200_LIBDIR"]).resolve() if os.environ.get("SILO_B200_LIBDIR") else PKG
SIGMA = {0: 16, 1: 25}

_host = None


def build(force=False):
    """Compile libsilo_b200.so (nvcc, sm_100a) and libsilo_host.so (g++) in-tree."""
    args = ["make", "-C", str(PKG)] + (["-B"] if force else []) + ["all"]
    subprocess.run(args, check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)


def host():
    global _host
    if _host is not None:
        return _host
    for name in ("libsilo_b200.so", "libsilo_host.so"):
        if not (LIBDIR / name).exists():
            raise RuntimeError(f"{name} is not built: run `make -C {PKG}` (there is no fallback path)")
    ctypes.CDLL(str(LIBDIR / "libsilo_b200.so"), mode=ctypes.RTLD_GLOBAL)
    lib = ctypes.CDLL(str(LIBDIR / "libsilo_host.so"), mode=ctypes.RTLD_GLOBAL)
    vp, cp, u32, u64, sz = ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_size_t
    pu32 = ctypes.POINTER(ctypes.c_uint32)
    sig = {
        "silo_last_error": (cp, []),
        "silo_free": (None, [vp]),
        "silo_db_build_from_dir": (vp, [cp]),
        "silo_db_build_from_dir_device": (vp, [cp, ctypes.c_int]),
        "silo_db_compare_stores": (ctypes.c_int64, [vp, vp]),
        "silo_db_free": (None, [vp]),
        "silo_db_partition_count": (u32, [vp]),
        "silo_db_partition_rows": (u32, [vp, u32]),
        "silo_db_info_json": (vp, [vp]),
        "silo_engine_create": (vp, [vp, ctypes.c_int, pu32, u32]),
        "silo_engine_free": (None, [vp]),
        "silo_engine_handle": (vp, [vp]),
        "silo_engine_execute": (vp, [vp, cp]),
        "silo_engine_mutation_counts": (ctypes.c_int, [vp, cp, cp, vp, sz]),
        "silo_engine_filter_ids": (ctypes.c_int, [vp, cp, u32, ctypes.POINTER(vp), ctypes.POINTER(sz)]),
        "silo_engine_eval_optree": (ctypes.c_int, [ctypes.c_int, cp, ctypes.POINTER(vp), ctypes.POINTER(sz)]),
        "silo_query_begin": (vp, [vp, cp]),
        "silo_query_free": (None, [vp]),
        "silo_query_sequence_count": (u32, [vp]),
        "silo_query_sequence_name": (cp, [vp, u32]),
        "silo_query_table_len": (u64, [vp, u32]),
        "silo_query_local_count": (ctypes.c_int, [vp, pu32]),
        "silo_query_local_table_device": (ctypes.c_int, [vp, u32, vp, u64, ctypes.POINTER(ctypes.c_int)]),
        "silo_query_set_table": (ctypes.c_int, [vp, u32, vp, u64]),
        "silo_query_set_table_device": (ctypes.c_int, [vp, u32, vp]),
        "silo_query_finish": (vp, [vp, u32]),
        "silo_rows_from_table": (vp, [vp, cp, cp, vp, u64]),
        "silo_prepared_create": (vp, [vp, cp, ctypes.c_int]),
        "silo_prepared_free": (None, [vp]),
        "silo_prepared_run_device": (ctypes.c_int, [vp, vp, ctypes.POINTER(u64)]),
        "silo_prepared_run": (vp, [vp, vp]),
        "silo_prepared_run_many": (ctypes.c_int, [vp, ctypes.POINTER(vp), u32, ctypes.POINTER(vp)]),
        "silo_prepared_table": (ctypes.c_int, [vp, u32, vp, u64]),
        "silo_prepared_bytes": (ctypes.c_int, [vp, vp, u32, ctypes.POINTER(u64), ctypes.POINTER(u64)]),
        "silo_prepared_elapsed_ms": (ctypes.c_float, [vp]),
        "silo_engine_set_plan_cache": (None, [vp, ctypes.c_int]),
        "silo_engine_execute_sharded": (vp, [vp, cp, ctypes.c_int]),
        "silo_engine_execute_partial": (vp, [vp, cp]),
        "silo_merge_partials": (vp, [vp, cp, ctypes.POINTER(cp), u32]),
        "silo_alias_apply": (vp, [cp, cp, ctypes.c_int]),
        "silo_parse_fasta": (vp, [cp, ctypes.c_size_t]),
        "silo_parse_database_config": (vp, [cp]),
        "silo_engine_filter_portable": (ctypes.c_int, [vp, cp, u32, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]),
        "silo_portable_ids": (ctypes.c_int, [cp, ctypes.c_size_t, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]),
        "silo_synth_generate": (vp, [cp, u64, u64, ctypes.c_int, ctypes.c_int, u32, u32, u32, u32, u32]),
        "silo_synth_owned_count": (u32, [vp]),
        "silo_synth_owned": (u32, [vp, u32]),
        "silo_synth_free": (None, [vp]),
        "silo_synth_db": (vp, [vp]),
        "silo_synth_clade_mutation_count": (u32, [vp]),
        "silo_synth_clade_mutation": (None, [vp, u32, pu32, ctypes.c_char_p]),
        "silo_synth_lineage_name": (cp, [vp, u32]),
        "silo_synth_verify_against_builder": (ctypes.c_int64, [vp]),
        "silo_synth_verify_against_builder_on": (ctypes.c_int64, [vp, ctypes.c_int]),
        "silo_test_worker_pool": (ctypes.c_int64, [u32, u32, u32, ctypes.c_int32]),
        "silo_literal_create": (vp, []),
        "silo_literal_free": (None, [vp]),
        "silo_literal_db": (vp, [vp]),
        "silo_literal_add_sequence": (ctypes.c_int, [vp, ctypes.c_int, cp, cp]),
        "silo_literal_add_partition": (ctypes.c_int, [vp, u32, pu32]),
        "silo_literal_set_bitmap": (ctypes.c_int, [vp, u32, cp, u32, ctypes.c_char, vp, u64, ctypes.c_int]),
        "silo_literal_set_flags": (ctypes.c_int, [vp, u32, cp, u32, ctypes.c_char, ctypes.c_char]),
        "silo_literal_set_missing": (ctypes.c_int, [vp, u32, cp, u32, vp, u32, ctypes.c_int]),
        "silo_literal_set_missing_bulk": (ctypes.c_int, [vp, u32, cp, u32, vp, vp, ctypes.c_int]),
        "silo_literal_add_values": (ctypes.c_int, [vp, u32, cp, u32, ctypes.c_char, vp, u64, u64, u64]),
        "silo_literal_flip_most_numerous": (ctypes.c_int, [vp, u32, cp, u32, u32, cp]),
        "silo_literal_delete_most_numerous": (ctypes.c_int, [vp, u32, cp, u32, u32, cp]),
        "silo_literal_get_bitmap": (ctypes.c_int, [vp, u32, cp, u32, ctypes.c_char, ctypes.POINTER(vp), ctypes.POINTER(sz), cp, cp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _host = lib
    return lib


class SiloError(RuntimeError):
    """code 1 = QueryParseException (HTTP 400 in the reference), 2 = compilation, 3 = anything else (500)."""

    def __init__(self, message):
        super().__init__(message)
        kind = message.split(":", 1)[0]
        self.kind = kind
        self.message = message.split(":", 1)[1] if ":" in message else message


def _error():
    return SiloError(host().silo_last_error().decode())


def _take_string(ptr):
    if not ptr:
        raise _error()
    try:
        return ctypes.string_at(ptr).decode()
    finally:
        host().silo_free(ptr)


def _take_ids(ptr, n):
    try:
        return np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(ctypes.c_uint32)), shape=(max(n, 1),))[:n].copy()
    finally:
        host().silo_free(ptr)


class Database:
    """A silo::Database built on the host (dataset directory or synthetic generator)."""

    def __init__(self, handle, owner=None):
        self.handle = handle
        self._owner = owner   # synthetic handle that owns the Database
        self.owned = None     # synthetic multi-GPU: partitions built on this rank

    @classmethod
    def from_directory(cls, path):
        h = host().silo_db_build_from_dir(str(path).encode())
        if not h:
            raise _error()
        return cls(h)

    @classmethod
    def from_directory_device_build(cls, path, device=0):
        """the same database with every sequence store built on the GPU (sb_index_build) instead of the host builder"""
        h = host().silo_db_build_from_dir_device(str(path).encode(), device)
        if not h:
            raise _error()
        return cls(h)

    def store_differences(self, other):
        """number of differing bitmaps / flags / missing-symbol bitmaps between the sequence stores of two databases"""
        d = host().silo_db_compare_stores(self.handle, other.handle)
        if d < 0:
            raise _error()
        return d

    @classmethod
    def synthetic(cls, n_sequences, seed, with_amino_acids=False, explicit_sequences=False, genome_length=0, n_threads=0,
                  reference_genomes=None, world=1, rank=0, row_slices=1):
        ref = reference_genomes or ROOT / "tests" / "golden" / "exampleDataset1000Sequences" / "reference_genomes.json"
        s = host().silo_synth_generate(str(ref).encode(), n_sequences, seed, int(with_amino_acids), int(explicit_sequences), genome_length, n_threads, world, rank, row_slices)
        if not s:
            raise _error()
        db = cls(host().silo_synth_db(s), owner=s)
        db.owned = [host().silo_synth_owned(s, i) for i in range(host().silo_synth_owned_count(s))]
        return db

    @property
    def n_partitions(self):
        return host().silo_db_partition_count(self.handle)

    def partition_rows(self, p):
        return host().silo_db_partition_rows(self.handle, p)

    def info(self):
        return json.loads(_take_string(host().silo_db_info_json(self.handle)))

    def clade_mutations(self):
        out = []
        if self._owner is None:
            return out
        pos, sym = ctypes.c_uint32(), ctypes.create_string_buffer(2)
        for i in range(host().silo_synth_clade_mutation_count(self._owner)):
            host().silo_synth_clade_mutation(self._owner, i, ctypes.byref(pos), sym)
            out.append((pos.value, sym.raw[:1].decode()))
        return out

    def lineage_names(self):
        return [host().silo_synth_lineage_name(self._owner, i).decode() for i in range(96)]

    def verify_against_builder(self, device=-1):
        """rebuilds every store from the explicit row strings -- on the host (device < 0) or with the device index build -- and
        counts the differences to the generated index"""
        d = host().silo_synth_verify_against_builder_on(self._owner, device)
        if d < 0:
            raise _error()
        return d

    def close(self):
        if self._owner is not None:
            host().silo_synth_free(self._owner)
        elif self.handle:
            host().silo_db_free(self.handle)
        self.handle = None
        self._owner = None


class LiteralDatabase(Database):
    """A Database built bitmap by bitmap (host/capi_literal.cpp): store states that no builder produces -- flipped symbols,
    flipped + deleted at one position, run containers in high keys, exact cardinalities and row counts. Position writers
    follow src/silo/storage/position.cpp:24-127."""

    def __init__(self, sequences):
        """sequences: {name: reference text} (nucleotide) or {name: (alphabet_id, reference text)}"""
        self._lit = host().silo_literal_create()
        for name, ref in sequences.items():
            alphabet, text = ref if isinstance(ref, tuple) else (0, ref)
            self._check(host().silo_literal_add_sequence(self._lit, alphabet, name.encode(), text.encode()))
        super().__init__(host().silo_literal_db(self._lit))
        self.owned = []

    @staticmethod
    def _check(rc):
        if rc != 0:
            raise _error()

    @staticmethod
    def _ids(ids):
        a = np.ascontiguousarray(ids, dtype=np.uint32)
        return a, a.ctypes.data

    def add_partition(self, sequence_count):
        index = ctypes.c_uint32()
        self._check(host().silo_literal_add_partition(self._lit, sequence_count, ctypes.byref(index)))
        self.owned.append(index.value)
        return index.value

    def set_bitmap(self, partition, name, pos, symbol, ids, run_optimize=True):
        a, ptr = self._ids(ids)
        self._check(host().silo_literal_set_bitmap(self._lit, partition, name.encode(), pos, symbol.encode(), ptr, a.size, int(run_optimize)))

    def set_flags(self, partition, name, pos, flipped=None, deleted=None):
        self._check(host().silo_literal_set_flags(self._lit, partition, name.encode(), pos, (flipped or "\0").encode(), (deleted or "\0").encode()))

    def set_missing(self, partition, name, row, ranges, run_optimize=True):
        a, ptr = self._ids(np.asarray(ranges, dtype=np.uint32).reshape(-1))
        self._check(host().silo_literal_set_missing(self._lit, partition, name.encode(), row, ptr, a.size // 2, int(run_optimize)))

    def set_missing_bulk(self, partition, name, range_off, ranges, run_optimize=True):
        """rows 0 .. len(range_off) - 2: row r misses the [start, end) position ranges ranges[range_off[r]:range_off[r + 1]]"""
        off = np.ascontiguousarray(range_off, dtype=np.uint64)
        a, ptr = self._ids(np.asarray(ranges, dtype=np.uint32).reshape(-1))
        self._check(host().silo_literal_set_missing_bulk(self._lit, partition, name.encode(), off.size - 1, off.ctypes.data, ptr, int(run_optimize)))

    def add_values(self, partition, name, pos, symbol, ids, offset, interval):
        a, ptr = self._ids(ids)
        self._check(host().silo_literal_add_values(self._lit, partition, name.encode(), pos, symbol.encode(), ptr, a.size, offset, interval))

    def flip_most_numerous(self, partition, name, pos, sequence_count):
        out = ctypes.create_string_buffer(2)
        self._check(host().silo_literal_flip_most_numerous(self._lit, partition, name.encode(), pos, sequence_count, out))
        return out.raw[:1].decode() if out.raw[0] else None

    def delete_most_numerous(self, partition, name, pos, sequence_count):
        out = ctypes.create_string_buffer(2)
        self._check(host().silo_literal_delete_most_numerous(self._lit, partition, name.encode(), pos, sequence_count, out))
        return out.raw[:1].decode() if out.raw[0] else None

    def get_bitmap(self, partition, name, pos, symbol):
        """(ids, flipped symbol or None, deleted symbol or None)"""
        ptr, n = ctypes.c_void_p(), ctypes.c_size_t()
        f, d = ctypes.create_string_buffer(2), ctypes.create_string_buffer(2)
        self._check(host().silo_literal_get_bitmap(self._lit, partition, name.encode(), pos, symbol.encode(), ctypes.byref(ptr), ctypes.byref(n), f, d))
        return _take_ids(ptr, n.value).tolist(), (f.raw[:1].decode() if f.raw[0] else None), (d.raw[:1].decode() if d.raw[0] else None)

    def close(self):
        if self._lit:
            host().silo_literal_free(self._lit)
        self._lit = None
        self.handle = None


class Engine:
    """GPU-resident index of (a subset of) the partitions of a Database + query execution."""

    def __init__(self, db, device=0, partitions=None):
        self.db = db
        parts = list(partitions) if partitions is not None else []
        arr = (ctypes.c_uint32 * max(len(parts), 1))(*parts)
        self.handle = host().silo_engine_create(db.handle, device, arr, len(parts))
        if not self.handle:
            raise _error()
        self.sb = host().silo_engine_handle(self.handle)

    def execute(self, query):
        q = query if isinstance(query, str) else json.dumps(query)
        return json.loads(_take_string(host().silo_engine_execute(self.handle, q.encode())))["queryResult"]

    def execute_text(self, query_json):
        """The response body as the reference's HTTP handler writes it ({"queryResult": [...]}), not parsed."""
        return _take_string(host().silo_engine_execute(self.handle, query_json.encode() if isinstance(query_json, str) else query_json))

    def set_plan_cache(self, enabled):
        """the engine's query-text -> prepared-query cache (on by default); switching clears it"""
        host().silo_engine_set_plan_cache(self.handle, int(enabled))

    def execute_sharded(self, query_json, is_root):
        """multi-GPU: every rank calls with the same text in the same order; rank 0 gets the response text"""
        return _take_string(host().silo_engine_execute_sharded(self.handle, query_json.encode() if isinstance(query_json, str) else query_json, int(is_root)))

    def execute_partial(self, query):
        """Aggregated (with or without groupByFields) on a sharded database: this rank's entries, unordered (see merge_partials)"""
        q = query if isinstance(query, str) else json.dumps(query)
        return _take_string(host().silo_engine_execute_partial(self.handle, q.encode()))

    def mutation_counts(self, query, sequence_name, table_len):
        q = query if isinstance(query, str) else json.dumps(query)
        out = np.zeros(table_len, dtype=np.uint32)
        rc = host().silo_engine_mutation_counts(self.handle, q.encode(), sequence_name.encode(), out.ctypes.data, table_len)
        if rc not in (0, 4):
            raise _error()
        return out, rc == 0

    def filter_ids(self, filter_expression, local_partition):
        f = filter_expression if isinstance(filter_expression, str) else json.dumps(filter_expression)
        ptr, n = ctypes.c_void_p(), ctypes.c_size_t()
        if host().silo_engine_filter_ids(self.handle, f.encode(), local_partition, ctypes.byref(ptr), ctypes.byref(n)) != 0:
            raise _error()
        return _take_ids(ptr, n.value)

    def filter_portable(self, filter_expression, local_partition):
        """the filter's rows as CRoaring portable bytes (sb_result_export_portable)"""
        f = filter_expression if isinstance(filter_expression, str) else json.dumps(filter_expression)
        ptr, n = ctypes.c_void_p(), ctypes.c_size_t()
        if host().silo_engine_filter_portable(self.handle, f.encode(), local_partition, ctypes.byref(ptr), ctypes.byref(n)) != 0:
            raise _error()
        try:
            return ctypes.string_at(ptr, n.value)
        finally:
            host().silo_free(ptr)

    def close(self):
        if self.handle:
            host().silo_engine_free(self.handle)
            self.handle = None


def eval_optree(tree, device=0):
    """Evaluates a literal operator tree (operator known-answer tests) on the GPU; returns sorted ids."""
    t = tree if isinstance(tree, str) else json.dumps(tree)
    ptr, n = ctypes.c_void_p(), ctypes.c_size_t()
    if host().silo_engine_eval_optree(device, t.encode(), ctypes.byref(ptr), ctypes.byref(n)) != 0:
        raise _error()
    return _take_ids(ptr, n.value)


class PoolStats(ctypes.Structure):
    _fields_ = [(n, ctypes.c_uint64) for n in (
        "array_containers", "array_values", "run_containers", "run_values", "run_count", "bitset_containers", "bitset_values",
        "directory_bytes", "payload_bytes", "missing_bytes", "full_table_bytes", "total_device_bytes",
        "n_stores", "n_partitions", "n_work_items", "n_chunks")]


def _sb():
    host()
    lib = ctypes.CDLL(str(LIBDIR / "libsilo_b200.so"))
    lib.sb_last_error.restype = ctypes.c_char_p
    for name in ("sb_timer_begin", "sb_timer_end", "sb_launch_count", "sb_k1_accumulated", "sb_pool_get_stats", "sb_transfer_bytes"):
        getattr(lib, name).restype = ctypes.c_int
    lib.sb_timer_begin.argtypes = [ctypes.c_void_p]
    lib.sb_timer_end.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_float)]
    lib.sb_launch_count.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64)]
    lib.sb_k1_accumulated.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_uint64)]
    lib.sb_transfer_bytes.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
    lib.sb_pool_get_stats.argtypes = [ctypes.c_void_p, ctypes.POINTER(PoolStats)]
    return lib


def _sb_check(rc):
    if rc != 0:
        raise SiloError("other:" + _sb().sb_last_error().decode())


def timer_begin(engine):
    _sb_check(_sb().sb_timer_begin(engine.sb))


def timer_end(engine):
    ms = ctypes.c_float()
    _sb_check(_sb().sb_timer_end(engine.sb, ctypes.byref(ms)))
    return ms.value


def launch_count(engine):
    n = ctypes.c_uint64()
    _sb_check(_sb().sb_launch_count(engine.sb, ctypes.byref(n)))
    return n.value


def transfer_bytes(engine):
    a, b = ctypes.c_uint64(), ctypes.c_uint64()
    _sb_check(_sb().sb_transfer_bytes(engine.sb, ctypes.byref(a), ctypes.byref(b)))
    return a.value, b.value


def k1_accumulated(engine, reset=False):
    ms, n = ctypes.c_double(), ctypes.c_uint64()
    _sb_check(_sb().sb_k1_accumulated(engine.sb, int(reset), ctypes.byref(ms), ctypes.byref(n)))
    return ms.value, n.value


def comm_unique_id():
    """rank 0: a fresh ncclUniqueId (128 bytes) to hand to every rank (sb_comm_unique_id)"""
    buf = ctypes.create_string_buffer(128)
    _sb_check(_sb().sb_comm_unique_id(buf))
    return buf.raw


def comm_init(engine, unique_id, rank, world):
    """every rank: from now on sb_mutations_device leaves the sum of the count tables of all ranks on rank 0 (NCCL reduce on the library's stream)"""
    lib = _sb()
    lib.sb_comm_init.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_int]
    _sb_check(lib.sb_comm_init(engine.sb, unique_id, rank, world))


def comm_free(engine):
    lib = _sb()
    lib.sb_comm_free.argtypes = [ctypes.c_void_p]
    _sb_check(lib.sb_comm_free(engine.sb))


def set_overlap(engine, enabled):
    lib = _sb()
    lib.sb_set_overlap.argtypes = [ctypes.c_void_p, ctypes.c_int]
    _sb_check(lib.sb_set_overlap(engine.sb, int(enabled)))


def pool_stats(engine):
    st = PoolStats()
    _sb_check(_sb().sb_pool_get_stats(engine.sb, ctypes.byref(st)))
    return {n: getattr(st, n) for n, _ in PoolStats._fields_}


class Prepared:
    """A query compiled, lowered, uploaded and located once (sb_query_prepare); run*() replays filter evaluation + action as
    one submission -- with use_graph as one CUDA graph launch."""

    def __init__(self, engine, query, use_graph=True):
        q = query if isinstance(query, str) else json.dumps(query)
        self.engine = engine
        self.handle = host().silo_prepared_create(engine.handle, q.encode(), int(use_graph))
        if not self.handle:
            raise _error()

    def run_device(self):
        """one device-resident step; returns the number of result rows (Aggregated: the count)"""
        n = ctypes.c_uint64()
        if host().silo_prepared_run_device(self.engine.handle, self.handle, ctypes.byref(n)) != 0:
            raise _error()
        return n.value

    def elapsed_ms(self):
        """CUDA-event time of the last run on the query's own stream (first kernel to last copy)"""
        return host().silo_prepared_elapsed_ms(self.handle)

    def run_text(self):
        return _take_string(host().silo_prepared_run(self.engine.handle, self.handle))

    def run(self):
        return json.loads(self.run_text())["queryResult"]

    def table(self, name_index, table_len):
        """uint32 count table of sequence `name_index` of the last run"""
        out = np.zeros(table_len, dtype=np.uint32)
        if host().silo_prepared_table(self.handle, name_index, out.ctypes.data, table_len) != 0:
            raise _error()
        return out

    def algorithmic_bytes(self, name_index=0):
        total, k1 = ctypes.c_uint64(), ctypes.c_uint64()
        if host().silo_prepared_bytes(self.engine.handle, self.handle, name_index, ctypes.byref(total), ctypes.byref(k1)) != 0:
            raise _error()
        return {"alg_bytes": total.value, "k1_bytes": k1.value}

    def close(self):
        if self.handle:
            host().silo_prepared_free(self.handle)
            self.handle = None


def merge_partials(db, query, partial_texts):
    """one rank: the response of an Aggregated query from the ranks' Engine.execute_partial texts (equal tuples add up, then order / limit)"""
    q = query if isinstance(query, str) else json.dumps(query)
    arr = (ctypes.c_char_p * max(len(partial_texts), 1))(*[t.encode() for t in partial_texts])
    return json.loads(_take_string(host().silo_merge_partials(db.handle, q.encode(), arr, len(partial_texts))))["queryResult"]


def alias_apply(alias_map, lineage, alias=False):
    """host PangoLineageAliasLookup (storage.hpp) built from {alias: [values]}: unalias(lineage), or alias(lineage) with alias=True"""
    return _take_string(host().silo_alias_apply(json.dumps(alias_map).encode(), lineage.encode(), 1 if alias else 0))


def parse_fasta(text):
    """[key, sequence] pairs of a FASTA text as the builder reads it (host/io_util.hpp readFastaEntries)"""
    data = text if isinstance(text, bytes) else text.encode()
    return json.loads(_take_string(host().silo_parse_fasta(data, len(data))))


def parse_database_config(text):
    """a database_config.yaml text as the builder understands it (host/builder.hpp parseDatabaseConfig)"""
    return json.loads(_take_string(host().silo_parse_database_config(text.encode() if isinstance(text, str) else text)))


def portable_ids(data):
    """the ids in CRoaring portable bytes, read by the host's restated reader (host/roaring.hpp)"""
    ptr, n = ctypes.c_void_p(), ctypes.c_size_t()
    if host().silo_portable_ids(data, len(data), ctypes.byref(ptr), ctypes.byref(n)) != 0:
        raise _error()
    return _take_ids(ptr, n.value)


def run_many(engine, prepared):
    """a batch of Prepared queries from this thread (sb_query_run_many): every query is enqueued before the first is waited for;
    returns their query results"""
    n = len(prepared)
    handles = (ctypes.c_void_p * max(n, 1))(*[p.handle for p in prepared])
    texts = (ctypes.c_void_p * max(n, 1))()
    if host().silo_prepared_run_many(engine.handle, handles, n, texts) != 0:
        raise _error()
    return [json.loads(_take_string(texts[i]))["queryResult"] for i in range(n)]


def rows_from_table(db, query, sequence_name, counts):
    """Root-rank result assembly from a reduced uint32 table in host memory (no device needed)."""
    q = query if isinstance(query, str) else json.dumps(query)
    counts = np.ascontiguousarray(counts, dtype=np.uint32)
    return json.loads(_take_string(host().silo_rows_from_table(db.handle, q.encode(), sequence_name.encode(), counts.ctypes.data, counts.size)))["queryResult"]


class DistributedQuery:
    """Multi-GPU protocol (one process per GPU): begin() evaluates the local filters, local tables are written
    into device memory for an NCCL reduce by the caller, finish() assembles the result on the root rank."""

    def __init__(self, engine, query):
        q = query if isinstance(query, str) else json.dumps(query)
        self.handle = host().silo_query_begin(engine.handle, q.encode())
        if not self.handle:
            raise _error()
        self.names = [host().silo_query_sequence_name(self.handle, i).decode() for i in range(host().silo_query_sequence_count(self.handle))]

    def table_len(self, i):
        return host().silo_query_table_len(self.handle, i)

    def local_count(self):
        c = ctypes.c_uint32()
        if host().silo_query_local_count(self.handle, ctypes.byref(c)) != 0:
            raise _error()
        return c.value

    def local_table_device(self, i, device_ptr, length):
        produced = ctypes.c_int()
        if host().silo_query_local_table_device(self.handle, i, device_ptr, length, ctypes.byref(produced)) != 0:
            raise _error()
        return produced.value != 0

    def set_table(self, i, host_array):
        if host().silo_query_set_table(self.handle, i, host_array.ctypes.data, host_array.size) != 0:
            raise _error()

    def set_table_device(self, i, device_ptr):
        if host().silo_query_set_table_device(self.handle, i, device_ptr) != 0:
            raise _error()

    def finish(self, reduced_count=0):
        return json.loads(_take_string(host().silo_query_finish(self.handle, reduced_count)))["queryResult"]

    def finish_text(self, reduced_count=0):
        """the response body, not parsed (what execute_text returns on one GPU)"""
        return _take_string(host().silo_query_finish(self.handle, reduced_count))

    def close(self):
        if self.handle:
            host().silo_query_free(self.handle)
            self.handle = None
