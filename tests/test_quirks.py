"""Compile rules the reference's fixtures do not pin, on a six-row database whose
answers are worked out by hand from the reference's source. The FilterExpression -> Operator compile code is host logic that the oracle
and the device path share (host/expressions.hpp), so comparing the two says nothing about it: these expectations do.

rows (ordered by key)    main (ref ACGT)   G1 (ref MK*)   G2 (ref M*)   pango_lineage
  0                      ACGT              MK*            M*            B.1
  1                      ACGA              MKK            M*            B.1.1
  2                      AC-T              LKK            M*            (null)
  3                      ANGT              MKK            M*            A
  4                      RCGT              XXX            XX            B.1
  5                      GCGT              BKK            M*            B.1.2
CPU: oracle. GPU: device path."""
import json

import pytest

BACKENDS = [pytest.param("oracle", id="oracle"), pytest.param("device", id="device", marks=pytest.mark.gpu)]
MAIN = ["ACGT", "ACGA", "AC-T", "ANGT", "RCGT", "GCGT"]
G1 = ["MK*", "MKK", "LKK", "MKK", "XXX", "BKK"]
G2 = ["M*", "M*", "M*", "M*", "XX", "M*"]
LINEAGE = ["B.1", "B.1.1", "", "A", "B.1", "B.1.2"]


@pytest.fixture(scope="module")
def quirk_db(libs, tmp_path_factory):
    d = tmp_path_factory.mktemp("quirks")
    (d / "database_config.yaml").write_text(
        "schema:\n  instanceName: quirks\n  metadata:\n    - name: key\n      type: string\n    - name: pango_lineage\n      type: pango_lineage\n  primaryKey: key\n")
    (d / "preprocessing_config.yaml").write_text('metadataFilename: "metadata.tsv"\nreferenceGenomeFilename: "reference_genomes.json"\n')
    (d / "reference_genomes.json").write_text(json.dumps({"nucleotideSequences": [{"name": "main", "sequence": "ACGT"}],
                                                          "genes": [{"name": "G1", "sequence": "MK*"}, {"name": "G2", "sequence": "M*"}]}))
    keys = [f"k{i}" for i in range(6)]
    (d / "metadata.tsv").write_text("key\tpango_lineage\n" + "".join(f"{k}\t{v}\n" for k, v in zip(keys, LINEAGE)))
    for name, rows in (("nuc_main", MAIN), ("gene_G1", G1), ("gene_G2", G2)):
        (d / f"{name}.fasta").write_text("".join(f">{k}\n{s}\n" for k, s in zip(keys, rows)))
    return libs[0].Database.from_directory(d)


@pytest.fixture(params=BACKENDS)
def ids(request, libs, quirk_db):
    pysilo, oracle = libs
    if request.param == "oracle":
        yield lambda flt: list(oracle.filter_ids(quirk_db, flt, 0))
    else:
        engine = pysilo.Engine(quirk_db, 0)
        yield lambda flt: list(engine.filter_ids(flt, 0))
        engine.close()


def nuc(pos, sym):
    return {"type": "NucleotideEquals", "position": pos, "symbol": sym}


def aa(gene, pos, sym):
    return {"type": "AminoAcidEquals", "sequenceName": gene, "position": pos, "symbol": sym}


def maybe(child):
    return {"type": "Maybe", "child": child}


def exact(child):
    return {"type": "Exact", "child": child}


def negate(child):
    return {"type": "Not", "child": child}


def test_rows_are_where_the_table_says(ids):
    assert ids({"type": "True"}) == [0, 1, 2, 3, 4, 5]
    for pos in range(1, 5):
        for sym in "ACGTRN-":
            assert ids(nuc(pos, sym)) == [r for r in range(6) if MAIN[r][pos - 1] == sym]


def test_q2_has_nucleotide_mutation(ids):
    """has_mutation.cpp:51-77: the list {A,C,G,T} "minus the reference" by std::remove without erase -- a reference T stays in"""
    has = lambda pos: {"type": "HasNucleotideMutation", "position": pos}
    assert ids(has(1)) == [5]                    # ref A: C, G, T, T -> only the G of row 5 (R of row 4 is ambiguous, not a mutation)
    assert ids(has(2)) == []                     # ref C: the N of row 3 is no mutation
    assert ids(has(3)) == []                     # ref G: the gap of row 2 is no mutation
    assert ids(has(4)) == [0, 1, 2, 3, 4, 5]     # ref T: {A,C,G,T} unchanged -> rows that carry the reference T match as well
    # under Maybe it is Not(Equals(reference)), compiled in mode NONE (:51-56): everything that is not exactly the reference
    assert ids(maybe(has(1))) == [4, 5]
    assert ids(maybe(has(2))) == [3]
    assert ids(maybe(has(3))) == [2]
    assert ids(maybe(has(4))) == [1]
    assert ids(exact(has(4))) == [0, 1, 2, 3, 4, 5]      # LOWER_BOUND takes the list path


def test_q3_has_amino_acid_mutation(ids):
    """has_aa_mutation.cpp:41-62: all symbols minus X minus the reference, same idiom -- a reference '*' (last of SYMBOLS) stays in"""
    has = lambda gene, pos: {"type": "HasAminoAcidMutation", "sequenceName": gene, "position": pos}
    assert ids(has("G1", 1)) == [2, 5]           # ref M: L of row 2, B of row 5 (B is in the list, only X is taken out)
    assert ids(has("G1", 2)) == []               # everybody K, row 4 X
    assert ids(has("G1", 3)) == [0, 1, 2, 3, 5]  # ref '*': the list keeps '*' -> every row that is not X, the '*' of row 0 included
    assert ids(maybe(has("G1", 1))) == [2, 4, 5]  # Not(Equals(M)): the X of row 4 counts now
    assert ids(maybe(has("G1", 3))) == [1, 2, 3, 4, 5]
    assert ids(has("G2", 1)) == []


def test_q4_equals_on_a_deleted_stop_codon_is_refused(libs, ids):
    """aa_symbol_equals.cpp:71-87: '*' deleted at G2:2 (the most numerous symbol there) -> the expansion contains itself; an error here"""
    pysilo, oracle = libs
    for flt in (aa("G2", 2, "*"), {"type": "HasAminoAcidMutation", "sequenceName": "G2", "position": 2}, negate(aa("G2", 2, "*"))):
        with pytest.raises((pysilo.SiloError, oracle.OracleError)) as err:
            ids(flt)
        assert "recurses without bound" in str(err.value)
    assert ids(aa("G1", 3, "*")) == [0]          # '*' is not the deleted symbol at G1:3 (K is)
    assert ids(aa("G1", 3, "K")) == [1, 2, 3, 5]  # the deleted symbol itself: And of the negations of all the others
    assert ids(aa("G2", 2, "X")) == [4]


def test_q5_the_dot_symbol(ids):
    """nucleotide_symbols.cpp:48-50: '.' is the gap for nucleotides; aa_symbol_equals.cpp:120-125: '.' is the reference symbol for amino acids"""
    assert ids(nuc(3, ".")) == [2] and ids(nuc(1, ".")) == []
    assert ids(aa("G1", 1, ".")) == [0, 1, 3]
    assert ids(aa("G1", 2, ".")) == [0, 1, 2, 3, 5]
    assert ids(aa("G1", 3, ".")) == [0]


def test_q9_ambiguity_modes(ids):
    """nucleotide_symbol_equals.cpp:28-73 (IUPAC supersets under UPPER_BOUND), negation.cpp:32 (Not swaps the mode), maybe / exact force it,
    nof.cpp:220-271 (exact N-Of under a mode = at least k and not at least k + 1, both in the SAME mode); amino acids ignore the mode"""
    assert ids(maybe(nuc(1, "A"))) == [0, 1, 2, 3, 4]          # A, R, M, W, D, H, V, N: the R of row 4
    assert ids(maybe(nuc(1, "G"))) == [4, 5]
    assert ids(maybe(nuc(2, "C"))) == [0, 1, 2, 3, 4, 5]       # N of row 3
    assert ids(maybe(nuc(3, "G"))) == [0, 1, 3, 4, 5]          # a gap is not ambiguous
    assert ids(exact(nuc(1, "A"))) == [0, 1, 2, 3]
    assert ids(negate(maybe(nuc(1, "A")))) == [5]
    assert ids(maybe(negate(nuc(1, "A")))) == [4, 5]           # Not swaps UPPER to LOWER: not exactly A
    assert ids(exact(negate(nuc(1, "A")))) == [5]              # Not swaps LOWER to UPPER: not possibly A
    assert ids(exact(maybe(nuc(1, "A")))) == [0, 1, 2, 3, 4]   # the inner one wins
    assert ids(maybe(exact(nuc(1, "A")))) == [0, 1, 2, 3]
    assert ids(maybe(aa("G1", 1, "M"))) == [0, 1, 3]           # neither B nor X widen an amino acid match
    children = [nuc(1, "A"), nuc(1, "G"), nuc(4, "A")]
    nof = lambda k, exactly: {"type": "N-Of", "numberOfMatchers": k, "matchExactly": exactly, "children": children}
    # matches per row in mode NONE: 1 2 1 1 0 1; under UPPER_BOUND: 1 2 1 1 2 1
    assert ids(nof(1, True)) == [0, 2, 3, 5] and ids(maybe(nof(1, True))) == [0, 2, 3, 5]
    assert ids(nof(2, True)) == [1] and ids(maybe(nof(2, True))) == [1, 4]
    assert ids(nof(1, False)) == [0, 1, 2, 3, 5] and ids(maybe(nof(1, False))) == [0, 1, 2, 3, 4, 5]
    assert ids(exact(nof(2, True))) == [1]
    assert ids(maybe(negate(nof(2, False)))) == [0, 2, 3, 4, 5]      # Not -> LOWER: at least two exact matches is row 1 only
    assert ids(negate(maybe(nof(2, False)))) == [0, 2, 3, 5]


def test_q13_lineage_filter(ids):
    """pango_lineage_filter.cpp:42-58: the query value is upper-cased; unknown lineages give no rows; the null lineage is stored as "" """
    lin = lambda value, sub: {"type": "PangoLineage", "column": "pango_lineage", "value": value, "includeSublineages": sub}
    assert ids(lin("b.1", False)) == [0, 4] and ids(lin("B.1", False)) == [0, 4]
    assert ids(lin("b.1", True)) == [0, 1, 4, 5]
    assert ids(lin("B", True)) == [0, 1, 4, 5] and ids(lin("B", False)) == []
    assert ids(lin("A", True)) == [3]
    assert ids(lin("B.1.3", True)) == [] and ids(lin("Q", False)) == []
    assert ids(negate(lin("B", True))) == [2, 3]
    assert ids(lin("", False)) == [2] and ids(lin("", True)) == [2]
    assert ids({"type": "PangoLineage", "column": "no_such_column", "value": "B.1", "includeSublineages": True}) == []      # :41-43 Empty, not an error
