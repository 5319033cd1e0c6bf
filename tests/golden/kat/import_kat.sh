#!/bin/bash
# Imports the known-answer cases of the reference's own end-to-end test table (scripts/test/simple_tests.pl:66-4424; it cannot run
# as shipped: Clone / Test::Deep / File::Which are not installed, and its CI builds binaries the fork no longer links) as DATA:
# the Perl array is evaluated by perl itself and dumped as JSON, then filtered to what this fork can do at all -- unpaired,
# end-to-end (no --local), sequence given inline or as FASTQ/FASTA text -- by tests/golden/kat/filter_kat.py, which also runs the
# UNMODIFIED reference binary (oracle/_ref) on every case and records whether the fork itself still meets the authors' expectation.
# Needs /root/reference; the product of this script (cases.json) is committed.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=${REF:-/root/reference}
sed -n '66,4424p' "$REF/scripts/test/simple_tests.pl" | sed 's/^my @cases = (/@cases = (/' > /tmp/kat_cases.pl
perl -MJSON::PP -e '
  my $compiled_with_sra = 0; my $should_test_bam = 0;
  local $/; open(my $f, "<", "/tmp/kat_cases.pl") or die; my $src = <$f>; close $f;
  my @cases; eval($src . "; 1") or die $@;
  print JSON::PP->new->canonical->pretty->encode(\@cases);' > /tmp/kat_all.json
python3 "$HERE/filter_kat.py" /tmp/kat_all.json "$HERE/cases.json"
