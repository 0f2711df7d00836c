"""Index build on the device (sb_index_build: SequenceStorePartition::fill + optimizeBitmaps, sequence_store.cpp:101-211)
against the host builder: every bitmap (set and serialized bytes), every deleted-symbol flag and every missing-symbol bitmap
must be identical -- on the reference's two datasets (all nucleotide segments and genes) and on synthetic rows."""
import pytest

from conftest import GOLDEN, golden_queries, load_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["exampleDataset", "exampleDataset1000Sequences"])
def test_device_built_stores_equal_host_built(libs, name, example_db, thousand_db):
    pysilo, _ = libs
    host_db = example_db if name == "exampleDataset" else thousand_db
    device_db = pysilo.Database.from_directory_device_build(GOLDEN / name, 0)
    try:
        assert device_db.n_partitions == host_db.n_partitions
        assert device_db.store_differences(host_db) == 0
    finally:
        device_db.close()


def test_device_build_on_synthetic_rows(libs, synth_small):
    # 3,000 rows x (3,000 nt + every gene), explicit row strings: null genomes, N runs at both ends and inside, ambiguity codes
    assert synth_small.verify_against_builder(device=-1) == 0
    assert synth_small.verify_against_builder(device=0) == 0


def test_queries_on_a_device_built_database(libs):
    pysilo, _ = libs
    db = pysilo.Database.from_directory_device_build(GOLDEN / "exampleDataset", 0)
    engine = pysilo.Engine(db, 0)
    try:
        for path in golden_queries()[::5]:
            case = load_case(path)
            assert engine.execute(case["query"]) == case["expectedQueryResult"], path.stem
    finally:
        engine.close()
        db.close()


def test_illegal_character_is_the_reference_s_error(libs, tmp_path):
    import shutil
    pysilo, _ = libs
    src = GOLDEN / "exampleDataset"
    dst = tmp_path / "broken"
    shutil.copytree(src, dst)
    fasta = dst / "nuc_testSecondSequence.fasta"
    text = fasta.read_text().splitlines()
    for i, line in enumerate(text):
        if line and not line.startswith(">"):
            text[i] = "J" + line[1:]
            break
    fasta.write_text("\n".join(text) + "\n")
    with pytest.raises(pysilo.SiloError, match="Illegal character"):
        pysilo.Database.from_directory_device_build(dst, 0)
