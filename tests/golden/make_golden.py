#!/usr/bin/env python3
"""Regenerates tests/golden/ from the reference's own test fixtures (run in the build container only;
/root/reference does not exist on the GPU box).

* copies the VCF/FASTA inputs of tests/test-files/{variant-graph,founder-sequences}/ verbatim (test DATA, not source);
* writes expected.json with the expectations transcribed from the reference's Catch2 tests:
    - tests/variant_graph.cc:249-338      per-node (node, ref_pos, aln_pos, ref_label, [(target, label)…]) + overlap reports
    - tests/founder_sequences.cc:118-188  cut_positions, assigned_samples (column-major init list, rows), exact A2M text
    - tests/transpose_matrix.cc:188-251   three single-bit known-answer cases
"""
import json
import os
import shutil

REF = "/root/reference/tests/test-files"
HERE = os.path.dirname(os.path.abspath(__file__))

for sub in ("variant-graph", "founder-sequences"):
    dst = os.path.join(HERE, sub)
    os.makedirs(dst, exist_ok=True)
    for f in sorted(os.listdir(os.path.join(REF, sub))):
        if f.endswith((".vcf", ".fa")):
            shutil.copyfile(os.path.join(REF, sub, f), os.path.join(dst, f))

g1 = [
    [0, 0, 0, "AAAA", []], [1, 4, 4, "A", [[2, "G"]]], [2, 5, 5, "A", []], [3, 6, 6, "A", [[4, "T"], [4, "CC"]]],
    [4, 7, 8, "A", [[5, "T"], [5, "GGGG"]]], [5, 8, 12, "A", [[7, "T"]]], [6, 9, 13, "A", [[8, "CC"]]],
    [7, 10, 14, "A", [[9, "GG"]]], [8, 11, 15, "A", []], [9, 12, 16, "AA", []], [10, 14, 18, "", []],
]
g3 = [
    [0, 0, 0, "T", [[10, "T"]]], [1, 1, 1, "GC", [[8, "C"]]], [2, 3, 3, "TG", [[3, "CCCC"]]], [3, 5, 7, "G", []],
    [4, 6, 8, "G", [[5, "T"]]], [5, 7, 9, "AG", []], [6, 9, 11, "GC", [[10, "TTTT"]]],
    [7, 11, 13, "A", [[8, "G"], [9, "C"]]], [8, 12, 14, "A", []], [9, 13, 15, "CC", []], [10, 15, 17, "", []],
]
g4 = [row[:] for row in g3]
g4[7] = [7, 11, 13, "A", [[8, "G"], [9, "C"], [9, ""]]]
g4[10] = [10, 15, 17, "GGGG", []]
g4.append([11, 19, 21, "", []])
overlap_1 = [{"sample": "SAMPLE2", "chrom_copy": 0, "ref_pos": 9, "var_id": "a5", "gt": 1}]

expected = {
    "variant_graph": [  # tests/variant_graph.cc:249-338
        {"vcf": "test-1a.vcf", "fa": "test-1.fa", "nodes": g1, "overlaps": overlap_1},
        {"vcf": "test-1b.vcf", "fa": "test-1.fa", "nodes": g1, "overlaps": overlap_1},
        {"vcf": "test-2.vcf", "fa": "test-2.fa", "overlaps": [], "nodes": [
            [0, 0, 0, "GC", [[4, "TTTT"]]], [1, 2, 2, "AA", [[2, "C"]]], [2, 4, 4, "C", [[3, "GG"]]], [3, 5, 6, "C", []], [4, 6, 7, "", []]]},
        {"vcf": "test-3.vcf", "fa": "test-3.fa", "nodes": g3, "overlaps": []},
        {"vcf": "test-4.vcf", "fa": "test-4.fa", "nodes": g4, "overlaps": []},
    ],
    "founder_sequences": [  # tests/founder_sequences.cc:118-188; founders=2, min_distance=0, keep_ref_edges=false
        {"vcf": "test-1.vcf", "fa": "test-1.fa", "cuts": [0, 1, 3, 5], "assigned_colmajor": [0, 6, 6, 3, 5, 8], "rows": 3,
         "a2m": ">REF\nCAA-AACTT-CCCGG-\n>1\nAAA-AACTT-CCAGG-\n>2\nCAA-AATTT-CCTGG-\n"},
        {"vcf": "test-1.vcf", "fa": "test-1-2.fa", "cuts": [0, 1, 3, 6], "assigned_colmajor": [0, 6, 6, 3, 5, 8], "rows": 3,
         "a2m": ">REF\nCAA-AACTT-CCCGG-AAAA\n>1\nAAA-AACTT-CCAGG-AAAA\n>2\nCAA-AATTT-CCTGG-AAAA\n"},
        {"vcf": "test-2.vcf", "fa": "test-2.fa", "cuts": [0, 3, 5], "assigned_colmajor": [6, 8, 0, 7], "rows": 2,
         "a2m": ">REF\nCAA-CTTCG-G\n>1\nCAA-CTTGG-G\n>2\nAAA-CTGGGGG\n"},
        {"vcf": "test-3.vcf", "fa": "test-3.fa", "cuts": [0, 1, 2, 3], "assigned_colmajor": [0, 6, 6, 3, 5, 8], "rows": 3,
         "a2m": ">REF\nCAA-CTT-CGG-\n>1\nAAA-CTT-AGG-\n>2\nCAA-TTT-TGG-\n"},
        {"vcf": "test-4.vcf", "fa": "test-4.fa", "cuts": [0, 2, 4, 6], "assigned_colmajor": [0, 6, 6, 3, 5, 8], "rows": 3,
         "a2m": ">REF\nTTTCAA-AACTT-CCCGG-\n>1\nTTTAAA-AACTT-CCAGG-\n>2\nTTTCAA-AATTT-CCTGG-\n"},
    ],
    "transpose_kat": [  # tests/transpose_matrix.cc:188-251: (rows, cols, set (row, col)) → transposed has (col, row)
        {"rows": 64, "cols": 128, "set": [1, 68]},
        {"rows": 128, "cols": 64, "set": [68, 1]},
        {"rows": 128, "cols": 128, "set": [68, 1]},
    ],
}
with open(os.path.join(HERE, "expected.json"), "w") as f:
    json.dump(expected, f, indent=1)
print("wrote", os.path.join(HERE, "expected.json"))
