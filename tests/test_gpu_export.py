"""sb_result_export_portable: a filter result as CRoaring portable bytes -- what the reference's host-side actions would turn back
into an OperatorResult with roaring::Roaring::readSafe (roaring_serialize.h:44). The bytes are read here twice: by a parser of the
public portable format written in this file, and by the host's restated reader; both must yield sb_result_export_ids' rows.
Array and bitset containers, several keys, the empty and the full set."""
import struct

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def parse_portable(data):
    """Roaring portable format without run containers: cookie 12346, container count, {key, cardinality - 1} x n, offsets x n, payloads"""
    cookie, n = struct.unpack_from("<II", data, 0)
    assert cookie == 12346
    out = []
    end = 8 + 8 * n
    for i in range(n):
        key, card_minus_1 = struct.unpack_from("<HH", data, 8 + 4 * i)
        (offset,) = struct.unpack_from("<I", data, 8 + 4 * n + 4 * i)
        assert offset == end, "containers follow each other without gaps"
        card = card_minus_1 + 1
        if card <= 4096:
            low = np.frombuffer(data, dtype="<u2", count=card, offset=offset).astype(np.uint32)
            assert np.all(low[1:] > low[:-1])
            end = offset + 2 * card
        else:
            words = np.frombuffer(data, dtype="<u8", count=1024, offset=offset)
            low = np.nonzero(np.unpackbits(words.view(np.uint8), bitorder="little"))[0].astype(np.uint32)
            assert len(low) == card
            end = offset + 8192
        out.append(low + (key << 16))
    assert end == len(data)
    return np.concatenate(out) if out else np.zeros(0, dtype=np.uint32)


@pytest.fixture(scope="module")
def engine(libs):
    pysilo, _ = libs
    db = pysilo.Database.synthetic(1_500_000, 0x5110000E, genome_length=1000)
    e = pysilo.Engine(db, 0)
    yield db, e
    e.close()
    db.close()


def test_portable_bytes_hold_the_rows_of_the_filter(libs, engine):
    pysilo, oracle = libs
    db, e = engine
    clade = db.clade_mutations()
    leaf = lambda k: {"type": "NucleotideEquals", "position": clade[k][0] + 1, "symbol": clade[k][1]}
    filters = [{"type": "True"}, {"type": "Not", "child": {"type": "True"}}, leaf(0), {"type": "Not", "child": leaf(0)},
               {"type": "N-Of", "numberOfMatchers": 2, "matchExactly": False, "children": [leaf(k) for k in range(0, 24, 3)]},
               {"type": "IntBetween", "column": "age", "from": 30, "to": 30}, {"type": "NucleotideEquals", "position": 700, "symbol": "N"}]
    largest = max(range(db.n_partitions), key=db.partition_rows)
    smallest = min(range(db.n_partitions), key=db.partition_rows)
    assert db.partition_rows(largest) > 65536
    arrays = bitsets = 0
    for p in (largest, smallest):
        for flt in filters:
            want = np.asarray(e.filter_ids(flt, p))
            data = e.filter_portable(flt, p)
            got = parse_portable(data)
            assert np.array_equal(got, want), (p, flt["type"])
            assert np.array_equal(np.asarray(pysilo.portable_ids(data)), want)
            assert np.array_equal(want, np.asarray(oracle.filter_ids(db, flt, p)))
            n = struct.unpack_from("<I", data, 4)[0]
            cards = [struct.unpack_from("<H", data, 10 + 4 * i)[0] + 1 for i in range(n)]
            arrays += sum(c <= 4096 for c in cards)
            bitsets += sum(c > 4096 for c in cards)
    assert arrays > 3 and bitsets > 3
