"""The `_probs.txt.gz` reader/writer (rpvg_b200/host/probs_io.cpp, format of rpvg:src/threaded_output_writer.cpp:42-95):
a fixture written by hand in the reference's layout, a round trip of a synthetic arena, and malformed input. Host code
only: runs without a GPU."""

import gzip
import os

import numpy as np
import pytest

from rpvg_b200 import probs_io, synth

HERE = os.path.dirname(os.path.abspath(__file__))


def test_reads_the_reference_layout():
    arena, names, lengths, eff = probs_io.read_probs(os.path.join(HERE, "golden", "example_probs.txt.gz"))
    assert arena.n_clusters == 2
    assert list(arena.read_offset) == [0, 2, 4] and list(arena.path_offset) == [0, 3, 4]
    assert names == ["ENST0001_H1", "ENST0001_H2", "ENST0002", "tx,with,commas"]
    assert list(lengths) == [1500, 1500, 820, 900]
    assert np.allclose(eff, [1312.4567, 1312.4567, 633.1, 711.5], rtol=0, atol=0)
    assert list(arena.row_ptr) == [0, 3, 4, 4, 5]          # third read class is noise-only
    assert list(arena.col) == [0, 1, 2, 2, 0]
    assert list(arena.prob) == [0.33329999, 0.33329999, 0.33340002, 0.95, 0.99]
    assert list(arena.noise) == [0.0001, 0.05, 1.0, 0.01]
    assert list(arena.count) == [12, 3, 1, 7]


def test_round_trip(tmp_path):
    arena = synth.transcripts_arena(seed=21, n_paths=400, n_reads=2500, n_max=60)
    rng = np.random.default_rng(3)
    names = [f"path_{i}" for i in range(arena.n_paths)]
    lengths = rng.integers(200, 5000, size=arena.n_paths).astype(np.uint32)
    eff = lengths - rng.uniform(0, 150, size=arena.n_paths)
    out = str(tmp_path / "dump_probs.txt.gz")
    probs_io.write_probs(out, arena, names, lengths, eff, prob_precision=1e-8)

    back, names2, lengths2, eff2 = probs_io.read_probs(out)
    # clusters without read classes are not written (as in the reference): compare the others, in order
    kept = [c for c in range(arena.n_clusters) if arena.read_offset[c + 1] > arena.read_offset[c]]
    ref = arena.select(kept)
    assert back.n_clusters == ref.n_clusters
    assert np.array_equal(back.read_offset, ref.read_offset) and np.array_equal(back.path_offset, ref.path_offset)
    assert np.array_equal(back.row_ptr, ref.row_ptr) and np.array_equal(back.col, ref.col) and np.array_equal(back.count, ref.count)
    assert np.allclose(back.prob, ref.prob, rtol=1e-7, atol=0) and np.allclose(back.noise, ref.noise, rtol=1e-7, atol=0)   # 8 digits
    kept_paths = np.concatenate([np.arange(arena.path_offset[c], arena.path_offset[c + 1]) for c in kept]).astype(int)
    assert names2 == [names[i] for i in kept_paths]
    assert np.array_equal(lengths2, lengths[kept_paths]) and np.allclose(eff2, eff[kept_paths], rtol=1e-7)

    # the text itself: header, path line, read-class lines with probability groups in ascending order
    with gzip.open(out, "rt") as fh:
        lines = fh.read().splitlines()
    assert lines[0] == "#" and lines[1].startswith("path_")
    first = lines[2].split(" ")
    assert int(first[0]) == ref.count[0] and all(":" in f for f in first[2:])
    probs = [float(f.split(":")[0]) for f in first[2:]]
    assert probs == sorted(probs)

    # writing what was read reproduces the file's text (the dump is a fixed point at 8 digits)
    again = str(tmp_path / "again_probs.txt.gz")
    probs_io.write_probs(again, back, names2, lengths2, eff2, prob_precision=1e-8)
    with gzip.open(again, "rt") as fh:
        assert fh.read().splitlines() == lines


@pytest.mark.parametrize("text, message", [
    ("name,1,1.0\n1 0.5\n", "before the first"),
    ("#\n", "ends after a cluster header"),
    ("#\nname,10\n", "name,length,effective_length"),
    ("#\na,10,9.5\n1 0.1 0.9:3\n", "out of range"),
    ("#\na,10,9.5\n1 0.1 0.9;0\n", "prob:ids"),
    ("#\na,10,9.5 b,10,9.5\n1 0.1 0.4:0 0.5:0\n", "listed twice"),
])
def test_malformed_input_is_rejected(tmp_path, text, message):
    path = str(tmp_path / "bad_probs.txt.gz")
    with gzip.open(path, "wt") as fh:
        fh.write(text)
    with pytest.raises(ValueError, match=message):
        probs_io.read_probs(path)


@pytest.mark.gpu
def test_dump_drives_the_gpu_path(ctx, tmp_path):
    """A dump is a workload source: dump -> arena -> EM on the device -> TPM, against the CPU oracle on the same arena."""
    from oracle import loader
    arena = synth.transcripts_arena(seed=23, n_paths=600, n_reads=4000, n_max=80)
    names = [f"p{i}" for i in range(arena.n_paths)]
    lengths = np.full(arena.n_paths, 1000, dtype=np.uint32)
    eff = np.linspace(300.0, 900.0, arena.n_paths)
    out = str(tmp_path / "run_probs.txt.gz")
    probs_io.write_probs(out, arena, names, lengths, eff)
    dumped, _, _, eff2 = probs_io.read_probs(out)

    abundances, noise_count, total_count, its = ctx.em_abundances(dumped)
    o_ab, o_noise, o_total, o_its, _ = loader.port_em_abundances(dumped)
    assert np.array_equal(total_count, o_total) and np.array_equal(its, o_its)
    nz = o_ab != 0
    assert np.array_equal(abundances != 0, nz)
    assert np.max(np.abs(abundances[nz] - o_ab[nz]) / o_ab[nz]) < 1e-6
    tpm, _ = ctx.tpm(abundances, eff2)
    assert abs(tpm.sum() - 1e6) < 1e-3
