"""Regenerates tests/golden/ from the read-only reference mount (run in the build container only).

Copies DATA, never source code:
  * exampleDataset/            the reference's 100-sequence test dataset inputs (testBaseData/exampleDataset)
  * exampleDataset1000Sequences/  BASELINE config C1 inputs
  * configs/                    the database_config files of config/database_config.test.cpp
  * ndjsonFiles/                the inputs of preprocessing/sequence_info.test.cpp
  * fastaFiles/                 the four inputs of common/fasta_reader.test.cpp (and their .zstdfasta twins)
  * fastaFilesWithMissingSequences/, ndjsonWithNullSequences/, ndjsonWithSqlKeywordField/, tsvWithSqlKeywordField/
                               the four tiny input directories of preprocessing/preprocessor.test.cpp (whole directories)
  * queries/*.json             the end-to-end golden vectors of endToEndTests/test/queries that are on the
                               hot path (Aggregated with and without groupByFields, Mutations, AminoAcidMutations)
  * invalidQueries/*.json      the invalid-query fixtures (expected HTTP 400 message)
  * queries_out_of_scope.txt   names of the fixtures that were skipped and why
"""
import json
import shutil
from pathlib import Path

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent

DATASET_FILES = {
    "exampleDataset": [
        "database_config.yaml", "preprocessing_config.yaml", "small_metadata_set.tsv", "reference_genomes.json",
        "pangolineage_alias.json", "nuc_main.fasta.zst", "nuc_testSecondSequence.fasta", "gene_S.fasta.xz",
    ] + [f"gene_{g}.fasta" for g in ["E", "M", "N", "ORF1a", "ORF1b", "ORF3a", "ORF6", "ORF7a", "ORF7b", "ORF8", "ORF9b"]],
    "exampleDataset1000Sequences": [
        "database_config.yaml", "preprocessing_config.yaml", "reference_genomes.json", "pangolineage_alias.json", "sample.ndjson.zst",
    ],
}
WHOLE_DATASETS = ["ndjsonFiles", "fastaFiles", "fastaFilesWithMissingSequences", "ndjsonWithNullSequences", "ndjsonWithSqlKeywordField", "tsvWithSqlKeywordField"]
HOT_ACTIONS = {"Mutations", "AminoAcidMutations", "Aggregated"}


def main():
    for name, files in DATASET_FILES.items():
        dst = OUT / name
        dst.mkdir(exist_ok=True)
        for f in files:
            shutil.copyfile(REF / "testBaseData" / name / f, dst / f)
    for name in WHOLE_DATASETS:
        dst = OUT / name
        dst.mkdir(exist_ok=True)
        for f in sorted((REF / "testBaseData" / name).iterdir()):
            shutil.copyfile(f, dst / f.name)
    (OUT / "configs").mkdir(exist_ok=True)
    for f in sorted((REF / "testBaseData").glob("test_database_config*.yaml")):      # config/database_config.test.cpp
        shutil.copyfile(f, OUT / "configs" / f.name)
    skipped = []
    for sub in ["queries", "invalidQueries"]:
        dst = OUT / sub
        if dst.exists():
            shutil.rmtree(dst)
        dst.mkdir()
        for path in sorted((REF / "endToEndTests" / "test" / sub).glob("*.json")):
            case = json.loads(path.read_text())
            action = case["query"].get("action", {}) if isinstance(case.get("query"), dict) else {}
            hot = action.get("type") in HOT_ACTIONS
            if sub == "queries" and not hot:
                skipped.append(f"{path.name}\t{action.get('type')}{' +groupByFields' if action.get('groupByFields') else ''}")
                continue
            shutil.copyfile(path, dst / path.name)
    (OUT / "queries_out_of_scope.txt").write_text("\n".join(skipped) + "\n")
    print("hot-path fixtures:", len(list((OUT / "queries").glob("*.json"))), "skipped:", len(skipped))


if __name__ == "__main__":
    main()
