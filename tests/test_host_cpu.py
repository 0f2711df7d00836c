"""CPU-only checks of the host side and of the C ABI library (no compute calls without a GPU)."""
import ctypes
import re

import numpy as np
import pytest

from conftest import ROOT


def test_abi_exports_every_declared_symbol(libs):
    header = (ROOT / "include" / "silo_b200.h").read_text()
    declared = set(re.findall(r"^(?:int|const char\*)\s+(sb_\w+)\(", header, re.M))
    assert len(declared) >= 18
    lib = ctypes.CDLL(str(ROOT / "lapis-silo_b200" / "libsilo_b200.so"))
    for name in sorted(declared):
        assert hasattr(lib, name), f"libsilo_b200.so does not export {name}"


def test_header_is_plain_c(tmp_path):
    """the drop-in boundary is a C ABI: include/silo_b200.h compiles as C99 with pedantic warnings as errors, no C++ in the signatures"""
    import subprocess
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    source = tmp_path / "abi.c"
    source.write_text('#include "silo_b200.h"\nint main(void) { sb_node node; sb_mutation_row row; sb_group_row group; (void)node; (void)row; (void)group; return 0; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-I", str(root / "include"), str(source)], check=True)


def test_integration_snippets_compile_and_link(libs, tmp_path):
    """the calls INTEGRATION.md shows (tests/integration_snippets.cpp) compile against the header with warnings as errors and link against
    libsilo_b200.so -- the document cannot drift from the ABI"""
    import subprocess
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    binary = tmp_path / "integration_snippets"
    subprocess.run(["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I", str(root / "include"), str(root / "tests" / "integration_snippets.cpp"), "-o", str(binary),
                    "-L", str(root / "lapis-silo_b200"), "-lsilo_b200", f"-Wl,-rpath,{root / 'lapis-silo_b200'}"], check=True)
    assert subprocess.run([str(binary)], check=True, capture_output=True, text=True).stdout.strip() == "links"


def test_no_cpu_fallback_without_gpu(libs):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    pysilo, _ = libs
    lib = ctypes.CDLL(str(ROOT / "lapis-silo_b200" / "libsilo_b200.so"))
    lib.sb_last_error.restype = ctypes.c_char_p
    db = ctypes.c_void_p()
    assert lib.sb_db_create(0, 1, ctypes.byref(db)) != 0          # fails loudly, never falls back
    assert lib.sb_last_error()
    with pytest.raises(pysilo.SiloError):
        pysilo.eval_optree({"op": "full", "row_count": 4})


def test_synthetic_generator_matches_reference_shaped_builder(synth_small):
    """generated index == index built from the explicit row strings by the restated
    SequenceStorePartition::fill / optimizeBitmaps (builder.hpp), bitmap by bitmap, byte by byte"""
    assert synth_small.n_partitions >= 20
    assert synth_small.verify_against_builder() == 0


def test_synthetic_is_deterministic(libs):
    pysilo, _ = libs
    a = pysilo.Database.synthetic(2000, 7, genome_length=2000).info()
    b = pysilo.Database.synthetic(2000, 7, genome_length=2000, n_threads=3).info()
    assert a == b


def test_oracle_fast_mode_equals_faithful_mode(libs, synth_small):
    _, oracle = libs
    query = {"action": {"type": "Mutations", "minProportion": 0.02},
             "filterExpression": {"type": "Not", "child": {"type": "PangoLineage", "column": "pangoLineage", "value": "B.1.3", "includeSublineages": True}}}
    slow, ok1 = oracle.mutation_counts(synth_small, query, "main", 16 * 3000, faithful=True)
    fast, ok2 = oracle.mutation_counts(synth_small, query, "main", 16 * 3000, faithful=False)
    assert ok1 and ok2 and slow.sum() > 0
    assert np.array_equal(slow, fast)
    aa = {"action": {"type": "AminoAcidMutations", "minProportion": 0.02}, "filterExpression": {"type": "True"}}
    assert oracle.execute(synth_small, aa, faithful=True) == oracle.execute(synth_small, aa, faithful=False)


def test_oracle_fast_mode_equals_faithful_mode_with_bitset_containers(libs):
    """The same on a partition that holds bitset containers and two container keys: the largest partition (~83,000 rows) of a
    1 M-row full-genome database (2,812 bitset containers in all), under a filter that keeps a few thousand of its rows so that
    the reference's O(|F| x L) contains() loop (mutations.cpp:76-82) stays in seconds. Fast mode is the authority at 10-40 M rows."""
    pysilo, oracle = libs
    db = pysilo.Database.synthetic(1_000_000, 0x51100002)
    try:
        stats = db.info()["sequences"]["main"]
        assert stats["numberOfBitsetContainers"] > 1000
        p = max(range(db.n_partitions), key=db.partition_rows)
        assert db.partition_rows(p) > 65536
        f = {"type": "And", "children": [{"type": "IntBetween", "column": "age", "from": 30, "to": 32},
                                         {"type": "Not", "child": {"type": "NucleotideEquals", "position": 12, "symbol": "N"}}]}
        n_f = len(oracle.filter_ids(db, f, p))
        assert 500 < n_f < 10000, n_f
        query = {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": f}
        slow, ok1 = oracle.mutation_counts_subset(db, query, "main", 16 * 29903, [p], faithful=True)
        fast, ok2 = oracle.mutation_counts_subset(db, query, "main", 16 * 29903, [p], faithful=False)
        assert ok1 and ok2 and slow.sum() > 29903 * 0.9 * n_f
        assert np.array_equal(slow, fast)
    finally:
        db.close()


def test_worker_pool_runs_every_job_and_propagates_errors(libs):
    """the host pool that shares per-partition compile + lower and the per-sequence tables of a query (host/engine.hpp)"""
    pysilo, _ = libs
    h = pysilo.host()
    for threads, jobs, rounds in ((6, 23, 50), (1, 1, 3), (3, 1000, 5), (6, 0, 2)):
        assert h.silo_test_worker_pool(threads, jobs, rounds, -1) == jobs * (jobs + 1) // 2
    assert h.silo_test_worker_pool(6, 23, 1, 7) == -1        # an exception inside a job comes back to the caller
    assert h.silo_test_worker_pool(6, 23, 3, -1) == 276      # and the pool is usable again afterwards (a fresh pool here)


def test_aggregated_partials_of_a_sharded_database_merge_to_the_whole(libs, synth_small):
    """host/group_by.hpp mergeAggregatedPartials (what rank 0 of a sharded database does): the unordered answers for two disjoint
    sets of partitions -- here from the oracle -- add up to the oracle's answer over all partitions, with ordering and limits"""
    import copy
    pysilo, oracle = libs
    clade = synth_small.clade_mutations()
    flt = {"type": "N-Of", "numberOfMatchers": 1, "matchExactly": False, "children": [{"type": "NucleotideEquals", "position": p + 1, "symbol": s} for p, s in clade[:5]]}
    parts = list(range(synth_small.n_partitions))
    shards = [parts[0::2], parts[1::2]]
    for action in [{"type": "Aggregated"}, {"type": "Aggregated", "groupByFields": ["region"], "orderByFields": ["region"]},
                   {"type": "Aggregated", "groupByFields": ["age", "pangoLineage", "qc", "date"], "orderByFields": [{"field": "count", "order": "descending"}, "pangoLineage", "age", "qc", "date"], "limit": 9, "offset": 2}]:
        query = {"action": action, "filterExpression": flt}
        partial = copy.deepcopy(query)
        for key in ("orderByFields", "limit", "offset"):
            partial["action"].pop(key, None)
        texts = [oracle.execute_subset_text(synth_small, partial, shard) for shard in shards]
        assert pysilo.merge_partials(synth_small, query, texts) == oracle.execute(synth_small, query)


def test_pango_lineage_alias_lookup_on_the_reference_test_parameters(libs):
    """host/storage.hpp PangoLineageAliasLookup against the parameter lists of the reference's own unit test
    (src/silo/storage/pango_lineage_alias.test.cpp:31-42 unaliasPangoLineage, :61-74 aliasPangoLineage, :76-86 the alias file of exampleDataset)"""
    pysilo, _ = libs
    resolve_map = {"X": ["A"], "XY": ["A.1"]}
    for lineage, want in [("", ""), ("SomeNotListedAlias", "SomeNotListedAlias"), ("X", "A"), ("XY", "A.1"), ("X.1.1", "A.1.1"), ("XYX.1.1", "XYX.1.1"), (".X", ".X")]:
        assert pysilo.alias_apply(resolve_map, lineage) == want
    unresolve_map = {"BA": ["B.1.1.529"], "BC": ["B.1.1.529.1.1.1"], "XA": ["B.1.1.7", "B.1.177"]}
    for lineage, want in [("", ""), ("SomeNotListedAlias", "SomeNotListedAlias"), ("B.1.1.529", "B.1.1.529"), ("B.1.1.529.1", "BA.1"), ("B.1.1.529.1.1", "BA.1.1"),
                          ("B.1.1.529.1.1.1", "BA.1.1.1"), ("B.1.1.529.1.1.1.1.1.1.1.1", "BC.1.1.1.1.1"), ("XA.1.1", "XA.1.1"), ("XA.1.1.1.1.1.1.1", "XA.1.1.1.1.1.1.1")]:
        assert pysilo.alias_apply(unresolve_map, lineage, alias=True) == want
    import json
    from pathlib import Path
    from_file = json.loads((Path(__file__).parent / "golden" / "exampleDataset" / "pangolineage_alias.json").read_text())
    for lineage, want in [("B", "B"), ("B.1", "B.1"), ("B.1.2", "B.1.2"), ("C", "B.1.1.1"), ("EP", "B.1.1.529.2.75.3.1.1.4")]:
        assert pysilo.alias_apply(from_file, lineage) == want


def test_query_error_messages_are_the_reference_s(libs, example_db):
    """the 111 validation texts of the reference's query parser (tests/golden/query_error_messages.json, extracted by make_messages.py):
    each is either a literal of the host's parser, or -- where the host composes the text from the expression's name -- produced by
    the query below that provokes it (QueryParseException -> "parse:<text>", HTTP 400 in the reference)"""
    import glob
    import json
    import re
    from pathlib import Path
    pysilo, oracle = libs
    root = Path(__file__).resolve().parent
    messages = json.loads((root / "golden" / "query_error_messages.json").read_text())
    host = re.sub(r'"\s*\n\s*"', "", "".join(Path(f).read_text() for f in glob.glob(str(root.parent / "lapis-silo_b200" / "host" / "*.hpp"))))
    provoked = {
        "The field 'child' is required in a Not expression": {"type": "Not"},
        "The field 'child' is required in a Maybe expression": {"type": "Maybe"},
        "The field 'child' is required in a Exact expression": {"type": "Exact"},
        "The field 'children' is required in an And expression": {"type": "And"},
        "The field 'children' in an And expression needs to be an array": {"type": "And", "children": {}},
        "The field 'children' is required in an Or expression": {"type": "Or"},
        "The field 'children' in an Or expression needs to be an array": {"type": "Or", "children": "x"},
        "The field 'children' is required in an N-Of expression": {"type": "N-Of", "numberOfMatchers": 1, "matchExactly": False},
        "The field 'children' in an N-Of expression needs to be an array": {"type": "N-Of", "numberOfMatchers": 1, "matchExactly": False, "children": 3},
        "The field 'position' is required in a NucleotideEquals expression": {"type": "NucleotideEquals", "symbol": "A"},
        "The field 'position' in a NucleotideEquals expression needs to be an unsigned integer greater than 0": {"type": "NucleotideEquals", "position": "1", "symbol": "A"},
        "The field 'position' is required in a AminoAcidEquals expression": {"type": "AminoAcidEquals", "sequenceName": "S", "symbol": "A"},
        "The field 'position' in a AminoAcidEquals expression needs to be an unsigned integer greater than 0": {"type": "AminoAcidEquals", "sequenceName": "S", "position": 0, "symbol": "A"},
        "The field 'position' is required in a HasNucleotideMutation expression": {"type": "HasNucleotideMutation"},
        "The field 'position' in a HasNucleotideMutation expression needs to be an unsigned integer": {"type": "HasNucleotideMutation", "position": -1},
        "The field 'position' is required in a HasAminoAcidMutation expression": {"type": "HasAminoAcidMutation", "sequenceName": "S"},
        "The field 'position' in a HasAminoAcidMutation expression needs to be an unsigned integer": {"type": "HasAminoAcidMutation", "sequenceName": "S", "position": 1.5},
        "The field 'value' in the InsertionContains expression does not contain a valid regex pattern": {"type": "InsertionContains", "column": "nucleotideInsertions", "position": 1, "value": "A(("},
    }
    assert len(messages) == 111
    literal = 0
    for m in messages:
        if m["text"] in host:
            literal += 1
            continue
        assert m["text"] in provoked, m
        with pytest.raises(oracle.OracleError) as err:
            oracle.execute(example_db, {"action": {"type": "Aggregated"}, "filterExpression": provoked[m["text"]]})
        assert err.value.kind == "parse" and err.value.message.startswith(m["text"]), (m, str(err.value))
        if "regex" not in m["text"]:
            assert err.value.message == m["text"]
    assert literal >= 90


def test_fasta_reader_on_the_reference_s_files(libs):
    """common/fasta_reader.test.cpp:9-84 on copies of its four input files: two entries; no newline at the end; a first line without '>';
    a key without a sequence line -- the last two are FastaFormatExceptions with the reference's texts (fasta_reader.cpp:19-21,41-43)"""
    from pathlib import Path
    pysilo, _ = libs
    files = Path(__file__).parent / "golden" / "fastaFiles"
    assert pysilo.parse_fasta((files / "test.fasta").read_bytes()) == [["Key1", "ACGT"], ["Key2", "CGTA"]]
    assert pysilo.parse_fasta((files / "no_end_new_line.fasta").read_bytes()) == [["Key", "ACGT"]]
    with pytest.raises(pysilo.SiloError) as err:
        pysilo.parse_fasta((files / "wrong_format.fasta").read_bytes())
    assert "Fasta key prefix '>' missing for key: ACGT" in str(err.value)
    with pytest.raises(pysilo.SiloError) as err:
        pysilo.parse_fasta((files / "missing_genome.fasta").read_bytes())
    assert "Missing genome sequence in line following key: Key_without_genome" in str(err.value)
    assert pysilo.parse_fasta(b"") == [] and pysilo.parse_fasta(b">k\n\n") == [["k", ""]]
    with pytest.raises(pysilo.SiloError):
        pysilo.parse_fasta(b">k\nACGT\n\n")      # an empty line where a key is expected


def test_reference_genomes_as_the_reference_reads_them(libs, example_db):
    """storage/reference_genomes.test.cpp:8-32, asked through the query path: the '.' symbol of AminoAcidEquals is the reference symbol
    (aa_symbol_equals.cpp:53-54) and NucleotideEquals / AminoAcidEquals refuse a position beyond the reference with its length in the text"""
    _, oracle = libs
    info = example_db.info()["sequences"]
    assert len([n for n in info if n in ("main", "testSecondSequence")]) == 2
    nuc = lambda name, pos, sym: {"type": "NucleotideEquals", "sequenceName": name, "position": pos, "symbol": sym}
    aa = lambda name, pos, sym: {"type": "AminoAcidEquals", "sequenceName": name, "position": pos, "symbol": sym}
    count = lambda flt: oracle.execute(example_db, {"action": {"type": "Aggregated"}, "filterExpression": flt})[0]["count"]
    for name, length, index, symbol in (("S", 1274, 3, "F"), ("ORF1a", 4401, 10, "K"), ("ORF9b", 98, 10, "A")):
        assert count(aa(name, index + 1, ".")) == count(aa(name, index + 1, symbol)) > 0
        for other in "ACDEFGHIKLMNPQRSTVWY":
            if other != symbol:
                assert count({"type": "And", "children": [aa(name, index + 1, "."), aa(name, index + 1, other)]}) == 0
        count(aa(name, length, "X"))      # the last position is there (its reference symbol is the stop codon: quirk Q4 forbids '.')
        with pytest.raises(oracle.OracleError) as err:
            count(aa(name, length + 1, "."))
        assert err.value.kind == "parse" and err.value.message == f"AminoAcidEquals position is out of bounds '{length + 1}' > '{length}'"
    for name, length, index, symbol in (("main", 29903, 0, "A"), ("testSecondSequence", 4, 1, "C")):
        # most rows carry the reference base: Mutations never reports it, and HasNucleotideMutation is the complement of it among A, C, G, T
        rows = oracle.execute(example_db, {"action": {"type": "Mutations", "sequenceName": name, "minProportion": 0.0}, "filterExpression": {"type": "True"}})
        assert all(r["mutation"][0] == symbol for r in rows if r["mutation"][1:-1] == str(index + 1))
        assert not any(r["mutation"] == f"{symbol}{index + 1}{symbol}" for r in rows)
        count(nuc(name, length, "A"))
        with pytest.raises(oracle.OracleError) as err:
            count(nuc(name, length + 1, "A"))
        assert err.value.kind == "parse" and err.value.message == f"NucleotideEquals position is out of bounds '{length + 1}' > '{length}'"


def test_database_config_reader_on_the_reference_s_files(libs):
    """config/database_config.test.cpp:52-176 on copies of its yaml files: column types per (type, generateIndex), the full test config,
    optional dateToSortBy / partitionBy, additional entries ignored, an unknown type and a missing metadata node refused"""
    from pathlib import Path
    pysilo, _ = libs
    configs = Path(__file__).parent / "golden" / "configs"
    read = lambda name: pysilo.parse_database_config((configs / name).read_text())
    config = read("test_database_config.yaml")
    assert config["instanceName"] == "sars_cov-2_minimal_test_config" and config["primaryKey"] == "gisaid_epi_isl"
    assert config["dateToSortBy"] == "date" and config["partitionBy"] == "pango_lineage"
    assert config["metadata"] == [["gisaid_epi_isl", "STRING"], ["date", "DATE"], ["unsorted_date", "DATE"], ["region", "INDEXED_STRING"], ["country", "INDEXED_STRING"],
                                  ["pango_lineage", "INDEXED_PANGOLINEAGE"], ["division", "INDEXED_STRING"], ["age", "INT"], ["qc_value", "FLOAT"],
                                  ["nucleotideInsertions", "NUC_INSERTION"], ["aminoAcidInsertions", "AA_INSERTION"]]
    assert read("test_database_config_without_date_to_sort_by.yaml")["dateToSortBy"] is None
    without_partition = read("test_database_config_without_partition_by.yaml")
    assert without_partition["partitionBy"] is None and without_partition["dateToSortBy"] == "date" and len(without_partition["metadata"]) == 10
    additional = read("test_database_config_with_additional_entries.yaml")
    assert [m[0] for m in additional["metadata"]] == ["gisaid_epi_isl", "date", "region", "country", "pangoLineage"]
    with pytest.raises(pysilo.SiloError) as err:
        read("test_database_config_with_invalid_metadata_type.yaml")
    assert "wrong_type" in str(err.value)
    with pytest.raises(pysilo.SiloError) as err:
        read("test_database_config_with_invalid_structure.yaml")
    assert 'invalid node; first invalid key: "metadata"' in str(err.value)
    # database_config.test.cpp:64-75: the column type of every (value type, generateIndex) pair
    typed = lambda kind, index: pysilo.parse_database_config(
        f"schema:\n  instanceName: x\n  metadata:\n    - name: testName\n      type: {kind}\n" + ("      generateIndex: true\n" if index else "") + "  primaryKey: testName\n")["metadata"][0][1]
    assert [typed("string", False), typed("string", True), typed("date", False), typed("pango_lineage", True), typed("int", False), typed("float", False),
            typed("insertion", False), typed("aaInsertion", False)] == ["STRING", "INDEXED_STRING", "DATE", "INDEXED_PANGOLINEAGE", "INT", "FLOAT", "NUC_INSERTION", "AA_INSERTION"]


def test_metadata_fields_of_the_config_must_be_in_the_input(libs, tmp_path):
    """preprocessing/metadata_info.test.cpp:8-31 (and :39-46 of metadata_info.cpp): a column of the database config that the metadata
    file (or the first ndjson record) does not have stops the build with the reference's text; the valid configs of :33-83 build"""
    import json
    import shutil
    from pathlib import Path
    pysilo, _ = libs
    golden = Path(__file__).parent / "golden"

    def directory(name, source, schema_columns, tail):
        d = tmp_path / name
        shutil.copytree(golden / source, d)
        (d / "database_config.yaml").write_text("schema:\n  instanceName: testInstanceName\n  metadata:\n" +
                                                "".join(f"    - name: {n}\n      type: {t}\n" for n, t in schema_columns) + tail)
        return d
    bad = [("gisaid_epi_isl", "string"), ("notInMetadata", "pango_lineage"), ("country", "string")]
    good = [("gisaid_epi_isl", "string"), ("pango_lineage", "pango_lineage"), ("date", "date"), ("country", "string")]
    with pytest.raises(pysilo.SiloError) as err:
        pysilo.Database.from_directory(directory("bad_tsv", "exampleDataset", bad, "  primaryKey: gisaid_epi_isl\n"))
    assert "The metadata field 'notInMetadata' which is contained in the database config is not contained in the input." in str(err.value)
    db = pysilo.Database.from_directory(directory("good_tsv", "exampleDataset", good, "  primaryKey: gisaid_epi_isl\n"))
    assert db.info()["sequenceCount"] == 100
    with pytest.raises(pysilo.SiloError) as err:
        pysilo.Database.from_directory(directory("bad_ndjson", "ndjsonWithSqlKeywordField", [("primaryKey", "string"), ("group", "string"), ("absent", "int")], "  primaryKey: primaryKey\n"))
    assert "The metadata field 'absent' which is contained in the database config is not contained in the input." in str(err.value)


def test_sequence_names_of_the_input_must_be_those_of_the_reference_genomes(libs, tmp_path):
    """preprocessing/sequence_info.test.cpp:12-65 on copies of its inputs: the 1,000-sequence ndjson validates against its own reference
    genomes; it fails against exampleDataset's (a second nucleotide sequence the input lacks); an input with a second nucleotide sequence
    or without the ORF genes fails against the 1,000-sequence references -- with the texts of sequence_info.cpp:107-150"""
    import shutil
    from pathlib import Path
    pysilo, _ = libs
    golden = Path(__file__).parent / "golden"
    thousand = golden / "exampleDataset1000Sequences"

    def directory(name, ndjson, references):
        d = tmp_path / name
        d.mkdir()
        for f in ("database_config.yaml", "pangolineage_alias.json"):
            shutil.copyfile(thousand / f, d / f)
        shutil.copyfile(ndjson, d / ndjson.name)
        shutil.copyfile(references, d / "reference_genomes.json")
        (d / "preprocessing_config.yaml").write_text(f'ndjsonInputFilename: "{ndjson.name}"\npangoLineageDefinitionFilename: "pangolineage_alias.json"\nreferenceGenomeFilename: "reference_genomes.json"\n')
        return d
    ok = pysilo.Database.from_directory(directory("ok", thousand / "sample.ndjson.zst", thousand / "reference_genomes.json"))
    assert ok.info()["sequenceCount"] == 1000
    with pytest.raises(pysilo.SiloError) as err:
        pysilo.Database.from_directory(directory("too_many_references", thousand / "sample.ndjson.zst", golden / "exampleDataset" / "reference_genomes.json"))
    assert "The aligned nucleotide sequence testSecondSequence which is contained in the reference sequences is not contained in the input file" in str(err.value)
    # oneline_second_nuc.json is not valid JSON (no comma before its second nucleotide sequence): refused by the parser here, by DuckDB's
    # in the reference (sequence_info.cpp:86-92); with the comma put in, the name check refuses it
    with pytest.raises(pysilo.SiloError) as err:
        pysilo.Database.from_directory(directory("too_many_in_json", golden / "ndjsonFiles" / "oneline_second_nuc.json.zst", thousand / "reference_genomes.json"))
    assert "JSON" in str(err.value)
    import ctypes
    zstd = ctypes.CDLL("libzstd.so.1")
    zstd.ZSTD_decompress.restype = ctypes.c_size_t
    zstd.ZSTD_decompress.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t]
    packed = (golden / "ndjsonFiles" / "oneline_second_nuc.json.zst").read_bytes()
    buffer = ctypes.create_string_buffer(1 << 20)
    n_bytes = zstd.ZSTD_decompress(buffer, len(buffer), packed, len(packed))
    text = buffer.raw[:n_bytes].decode()
    assert '"\n    "testSecondSequence"' in text
    repaired = tmp_path / "oneline_second_nuc_repaired.json"
    repaired.write_text(text.replace('"\n    "testSecondSequence"', '",\n    "testSecondSequence"'))
    with pytest.raises(pysilo.SiloError) as err:
        pysilo.Database.from_directory(directory("too_many_in_json_repaired", repaired, thousand / "reference_genomes.json"))
    assert "The aligned nucleotide sequence testSecondSequence which is contained in the input file" in str(err.value) and "is not contained in the reference sequences." in str(err.value)
    with pytest.raises(pysilo.SiloError) as err:
        pysilo.Database.from_directory(directory("too_few_genes", golden / "ndjsonFiles" / "oneline_without_ORF.json.zst", thousand / "reference_genomes.json"))
    assert "The aligned amino acid sequence ORF" in str(err.value) and "which is contained in the reference sequences is not contained in the input file" in str(err.value)
