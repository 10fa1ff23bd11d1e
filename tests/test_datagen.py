"""The two exact generators of lime_b200/datagen.py agree: the numpy prefix-doubling suffix sort used for the
config 1 stand-in (tests/golden/make_config1.py) against the naive sort of Python strings."""
import numpy as np
import pytest

from lime_b200 import datagen


@pytest.mark.parametrize("term", [b"$", b"\0", b"#"])
def test_fast_exact_generator_equals_the_naive_one(term):
    rng = np.random.default_rng(3)
    genomes = [datagen.random_dna(rng, 700), datagen.random_dna(rng, 500, "ACGTN"), "ACGT" * 40, "A" * 90]
    reads, _ = datagen.sample_reads(rng, genomes, 80, 40, err=0.05, iupac=0.02)
    reads += ["", "A", reads[0], reads[0], datagen.revcomp(reads[1])]
    a = datagen.exact_collection(reads, genomes, term)
    b = datagen.exact_collection_fast(reads, genomes, term)
    for x, y in zip(a, b):
        assert x.dtype == y.dtype and np.array_equal(x, y)
