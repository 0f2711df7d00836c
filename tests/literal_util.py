"""Hand-made SequenceStorePartitions for the parity tests: a symbol matrix (rows x positions) becomes a literal store in
a chosen per-position state, and the same matrix gives an independent expected count table (a plain histogram)."""
import numpy as np

NUC = "-ACGTRYSWKMBDHVN"      # enum order of include/silo/common/nucleotide_symbols.h:17-34
CODE = {c: i for i, c in enumerate(NUC)}
MISSING = CODE["N"]


def random_matrix(rng, n_rows, length, reference, p_mut=0.08, p_missing_edge=0.3, p_gap=0.02):
    """rows x positions matrix of symbol codes: mostly the reference symbol, point mutations, gaps, ambiguous codes,
    N at the ends of some rows and in a few inner stretches."""
    ref = np.array([CODE[c] for c in reference], dtype=np.uint8)
    m = np.broadcast_to(ref, (n_rows, length)).copy()
    mut = rng.random((n_rows, length)) < p_mut
    m[mut] = rng.integers(1, 5, size=int(mut.sum()), dtype=np.uint8)       # A C G T (may equal the reference: no mutation)
    gap = rng.random((n_rows, length)) < p_gap
    m[gap] = CODE["-"]
    amb = rng.random((n_rows, length)) < 0.003
    m[amb] = rng.integers(5, 15, size=int(amb.sum()), dtype=np.uint8)      # R .. V
    edge = rng.random(n_rows) < p_missing_edge
    head = np.where(edge, rng.integers(0, max(2, length // 8), n_rows), 0)
    tail = np.where(edge, rng.integers(0, max(2, length // 8), n_rows), 0)
    cols = np.arange(length)
    m[cols[None, :] < head[:, None]] = MISSING
    m[cols[None, :] >= (length - tail)[:, None]] = MISSING
    inner = np.nonzero(rng.random(n_rows) < 0.02)[0]
    for r in inner[:2000]:
        a = int(rng.integers(0, length - 1))
        m[r, a:a + int(rng.integers(1, 12))] = MISSING
    return m


def missing_ranges(matrix):
    """per row the [start, end) runs of N -> (range_off[n + 1] uint64, ranges[k][2] uint32)"""
    n, length = matrix.shape
    is_n = np.zeros((n, length + 2), dtype=np.int8)
    is_n[:, 1:-1] = matrix == MISSING
    d = np.diff(is_n, axis=1)
    starts = np.argwhere(d == 1)          # (row, start)
    ends = np.argwhere(d == -1)           # (row, end) in the same order
    ranges = np.stack([starts[:, 1], ends[:, 1]], axis=1).astype(np.uint32)
    counts = np.bincount(starts[:, 0], minlength=n)
    off = np.zeros(n + 1, dtype=np.uint64)
    off[1:] = np.cumsum(counts)
    return off, ranges


def fill_partition(db, partition, name, matrix, modes, run_optimize=True):
    """Stores the matrix as partition `partition` of sequence `name`. modes[pos] in
         "raw"          every symbol's bitmap stored as is, no flag
         "delete"       most numerous symbol emptied and recorded as deleted (what preprocessing leaves, position.cpp:102-127)
         "flip"         most numerous symbol stored complemented and recorded as flipped (position.cpp:71-100)
         "flip+delete"  most numerous deleted AND the runner-up stored complemented / flipped
    The most numerous symbol is the first in enum order with the strictly highest count (position.cpp:42-69)."""
    n, length = matrix.shape
    all_rows = np.arange(n, dtype=np.uint32)
    for pos in range(length):
        col = matrix[:, pos]
        counts = np.bincount(col, minlength=16)[:16].copy()
        counts[MISSING] = 0
        mode = modes[pos]
        order = np.argsort(-counts, kind="stable")      # stable: first in enum order on ties
        top = int(order[0]) if counts[order[0]] > 0 else None
        second = int(order[1]) if counts[order[1]] > 0 else None
        flipped = deleted = None
        if mode in ("delete", "flip+delete") and top is not None:
            deleted = top
        if mode == "flip" and top is not None:
            flipped = top
        if mode == "flip+delete" and second is not None:
            flipped = second
        for sym in range(16):
            if sym == MISSING or sym == deleted:
                continue
            if counts[sym] == 0 and sym != flipped:
                continue
            ids = np.nonzero(col == sym)[0].astype(np.uint32)
            if sym == flipped:
                ids = np.setdiff1d(all_rows, ids, assume_unique=True)
            db.set_bitmap(partition, name, pos, NUC[sym], ids, run_optimize)
        db.set_flags(partition, name, pos, NUC[flipped] if flipped is not None else None, NUC[deleted] if deleted is not None else None)
    off, ranges = missing_ranges(matrix)
    db.set_missing_bulk(partition, name, off, ranges)


def histogram_table(matrices, masks):
    """expected uint32[16][L] Mutations table: rows selected by masks[p] of partition matrices[p], N never counted"""
    length = matrices[0].shape[1]
    table = np.zeros((16, length), dtype=np.uint32)
    for m, mask in zip(matrices, masks):
        sel = m[mask]
        for sym in range(16):
            if sym != MISSING:
                table[sym] += (sel == sym).sum(axis=0).astype(np.uint32)
    return table.reshape(-1)
