import json
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "lapis-silo_b200"))
sys.path.insert(0, str(ROOT / "tests"))
sys.path.insert(0, str(ROOT))
GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def libs():
    """Builds the product libraries and the oracle if they are not there yet (in-tree, no install)."""
    if not (ROOT / "lapis-silo_b200" / "libsilo_host.so").exists() or not (ROOT / "lapis-silo_b200" / "libsilo_b200.so").exists():
        subprocess.run(["make", "-C", str(ROOT / "lapis-silo_b200"), "all"], check=True)
    if not (ROOT / "oracle" / "liboracle.so").exists():
        subprocess.run(["make", "-C", str(ROOT / "oracle")], check=True)
    import pysilo
    import oracle_bind
    pysilo.host()
    oracle_bind.lib()
    return pysilo, oracle_bind


@pytest.fixture(scope="session")
def example_db(libs):
    pysilo, _ = libs
    return pysilo.Database.from_directory(GOLDEN / "exampleDataset")


@pytest.fixture(scope="session")
def thousand_db(libs):
    pysilo, _ = libs
    return pysilo.Database.from_directory(GOLDEN / "exampleDataset1000Sequences")


@pytest.fixture(scope="session")
def synth_small(libs):
    """3,000 synthetic sequences over a 3,000 nt prefix + all genes, with explicit row strings."""
    pysilo, _ = libs
    return pysilo.Database.synthetic(3000, 0x51100003, with_amino_acids=True, explicit_sequences=True, genome_length=3000)


@pytest.fixture(scope="session")
def synth_medium(libs):
    """120,000 synthetic sequences, full genome, 23 partitions of at most ~10,000 rows: one container key, one slab, no bitset
    container (several keys / slabs / bitsets: tests/test_gpu_fullsize.py and tests/test_gpu_literal.py)."""
    pysilo, _ = libs
    return pysilo.Database.synthetic(120000, 0x51100002)


def golden_queries():
    return sorted((GOLDEN / "queries").glob("*.json"))


def load_case(path):
    return json.loads(Path(path).read_text())
