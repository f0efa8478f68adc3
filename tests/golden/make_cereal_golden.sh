#!/bin/bash
# Regenerates tests/golden/cereal/*.graph / *.cuts with the real cereal headers of the reference checkout
# (oracle/_ref/cereal_golden, built by `make -C oracle` when /root/reference is present). Run from the repository root.
set -e
make -C oracle _ref/cereal_golden
G=tests/golden
for t in founder-sequences/test-1 founder-sequences/test-2 founder-sequences/test-3 founder-sequences/test-4 variant-graph/test-2 variant-graph/test-3 variant-graph/test-4; do
	name=$(echo $t | tr / _)
	oracle/_ref/cereal_golden $G/$t.fa $G/$t.vcf 0 $G/cereal/$name.graph $G/cereal/$name.cuts
done
# a second cut-position file with a non-zero minimum distance
oracle/_ref/cereal_golden $G/founder-sequences/test-1.fa $G/founder-sequences/test-1.vcf 3 /tmp/v2m_unused.graph $G/cereal/founder-sequences_test-1.md3.cuts
ls -la $G/cereal
