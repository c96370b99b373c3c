"""Freezes the path / transcript / haplotype structure of the reference's example pantranscriptome
(/root/reference/example/pantranscriptome.txt.gz: Name, Length, Transcript, Haplotypes per path) into
tests/golden/c1_structure.npz, the fixture rpvg_b200.synth.example_structure() reads. The reference tree does not travel
to the GPU box; this small derived file does. Run where /root/reference exists:

    python tests/golden/make_c1_structure.py
"""

import gzip
import os

import numpy as np

SRC = os.path.join(os.environ.get("RPVG_ROOT", "/root/reference"), "example", "pantranscriptome.txt.gz")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "c1_structure.npz")


def main():
    transcripts, haplotypes = {}, {}
    path_transcript, path_length, hap_ptr, hap_ids = [], [], [0], []
    with gzip.open(SRC, "rt") as fh:
        header = fh.readline().rstrip("\n").split("\t")
        assert header == ["Name", "Length", "Transcript", "Haplotypes"]
        for line in fh:
            name, length, transcript, haps = line.rstrip("\n").split("\t")
            path_transcript.append(transcripts.setdefault(transcript, len(transcripts)))
            path_length.append(int(length))
            ids = sorted({haplotypes.setdefault(h, len(haplotypes)) for h in haps.split(",") if h})
            hap_ids.extend(ids)
            hap_ptr.append(len(hap_ids))
    assert len(haplotypes) < 65536
    np.savez_compressed(OUT, path_transcript=np.array(path_transcript, dtype=np.uint32), path_length=np.array(path_length, dtype=np.uint32),
                        hap_ptr=np.array(hap_ptr, dtype=np.uint32), hap_ids=np.array(hap_ids, dtype=np.uint16))
    print(f"{len(path_transcript)} paths, {len(transcripts)} transcripts, {len(haplotypes)} haplotype ids, {len(hap_ids)} (path, haplotype) pairs -> {OUT}",
          os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
