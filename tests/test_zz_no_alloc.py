"""sb_mutation_rows_from_table_static: the row selection with its scratch in a file-scope device array (no allocation
in the call). Collected last (file name): compares the new entry point with sb_mutation_rows_from_table through the C ABI,
on two calls in a row that share the array -- the second with fewer rows than the first left behind."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

VALID = [0, 1, 2, 3, 4]      # -, A, C, G, T (SymbolType::VALID_MUTATION_SYMBOLS of the nucleotide alphabet)
GENOME_LENGTH = 29903


class Row(ctypes.Structure):
    _fields_ = [("position", ctypes.c_uint32), ("symbol", ctypes.c_uint32), ("count", ctypes.c_uint32), ("total", ctypes.c_uint32)]


def _abi(pysilo):
    lib = pysilo._sb()
    args = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_double, ctypes.POINTER(ctypes.c_uint8), ctypes.c_uint32,
            ctypes.POINTER(Row), ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]
    for name in ("sb_mutation_rows_from_table", "sb_mutation_rows_from_table_static"):
        getattr(lib, name).restype = ctypes.c_int
        getattr(lib, name).argtypes = args
    return lib


def _rows(lib, fn_name, engine, device_ptr, min_proportion, valid, capacity):
    rows = (Row * capacity)()
    n = ctypes.c_size_t()
    v = (ctypes.c_uint8 * len(valid))(*valid)
    rc = getattr(lib, fn_name)(engine.sb, b"main", device_ptr, min_proportion, v, len(valid), rows, capacity, ctypes.byref(n))
    if rc != 0:
        raise RuntimeError(lib.sb_last_error().decode())
    assert n.value <= capacity
    return [(r.position, r.symbol, r.count, r.total) for r in rows[:n.value]]


def test_static_scratch_row_selection_equals_the_allocating_one(libs, example_db):
    import torch
    pysilo, _ = libs
    lib = _abi(pysilo)
    engine = pysilo.Engine(example_db, 0)
    try:
        table_len = 16 * GENOME_LENGTH
        tables = []
        for flt in ({"type": "True"}, {"type": "PangoLineage", "column": "pango_lineage", "value": "B.1.1.7", "includeSublineages": True}):
            counts, produced = engine.mutation_counts({"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": flt}, "main", table_len)
            assert produced
            tables.append(torch.from_numpy(counts.view(np.int32)).to("cuda:0"))
        torch.cuda.synchronize()
        capacity = GENOME_LENGTH * len(VALID)
        # two calls in a row on the shared array: many rows first (min_proportion 0), then few -- and a second table
        calls = [(tables[0], 0.0), (tables[0], 0.05), (tables[1], 0.5), (tables[1], 0.0)]
        sizes = []
        for table, min_proportion in calls:
            want = _rows(lib, "sb_mutation_rows_from_table", engine, table.data_ptr(), min_proportion, VALID, capacity)
            got = _rows(lib, "sb_mutation_rows_from_table_static", engine, table.data_ptr(), min_proportion, VALID, capacity)
            assert got == want      # integers: exact, same order
            sizes.append(len(got))
        assert sizes[0] > sizes[1] > 0 and sizes[3] > sizes[2] > 0
        # *n_rows beyond capacity: the count comes back, no rows are written (same contract as the allocating call)
        rows = (Row * 4)()
        n = ctypes.c_size_t()
        v = (ctypes.c_uint8 * len(VALID))(*VALID)
        assert lib.sb_mutation_rows_from_table_static(engine.sb, b"main", tables[0].data_ptr(), 0.0, v, len(VALID), rows, 4, ctypes.byref(n)) == 0
        assert n.value == sizes[0]
        # beyond the static array: a clear error, nothing launched
        with pytest.raises(RuntimeError, match="exceed the static scratch"):
            _rows(lib, "sb_mutation_rows_from_table_static", engine, tables[0].data_ptr(), 0.0, list(range(16)), 16)
    finally:
        engine.close()
