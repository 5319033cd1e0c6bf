#!/bin/bash
# Regenerates the golden vectors of this directory by running the UNMODIFIED reference
# (oracle/_ref/bowtie2-align-l, built by oracle/build_ref.sh from /root/reference) here.
# Inputs: the reference's shipped example (example/reference/lambda_virus.fa, first 1000
# records of example/reads/reads_1.fq -- BASELINE config 1) and seeded synthetic sets
# from tools/gen_data.py.  Output order depends on the thread schedule, so records are
# stored sorted.  Needs /root/reference; not run on the GPU box.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"; ROOT="$HERE/../.."; REF=${REF:-/root/reference}
B="$ROOT/oracle/_ref"; T=$(mktemp -d)
cp "$REF/example/reference/lambda_virus.fa" "$HERE/lambda_virus.fa"
head -4000 "$REF/example/reads/reads_1.fq" > "$HERE/lambda_reads_1k.fq"
run() { # name index reads opts
  OMP_NUM_THREADS=4 "$B/bowtie2-align-l" -x "$2" -U "$3" -S "$T/o.sam" -p 8 $4 2>/dev/null
  grep -v '^@' "$T/o.sam" | LC_ALL=C sort > "$HERE/$1.sam"
}
"$B/bowtie2-build-l" "$HERE/lambda_virus.fa" "$T/lambda" >/dev/null 2>&1
run lambda_default "$T/lambda" "$HERE/lambda_reads_1k.fq" ""
cp "$REF/example/reads/reads_1.fq" "$HERE/lambda_reads_10k.fq"     # BASELINE configs[0] in full
run lambda_default_10k "$T/lambda" "$HERE/lambda_reads_10k.fq" ""
run lambda_k3_vs   "$T/lambda" "$HERE/lambda_reads_1k.fq" "-k 3 --very-sensitive"
# synthetic: 300 kbp i.i.d. reference, 3 sequences, 20% repeat families
python "$ROOT/tools/gen_data.py" ref "$T/syn.fa" 300000 --seed 4242 --repeats 0.2 --nseq 3
"$B/bowtie2-build-l" "$T/syn.fa" "$T/syn" >/dev/null 2>&1
python "$ROOT/tools/gen_data.py" reads "$T/syn.fa" "$T/syn150.fq" 1500 150 --seed 7 --randqual --nfrac 0.002
python "$ROOT/tools/gen_data.py" reads "$T/syn.fa" "$T/syn100.fq" 1500 100 --seed 8 --subst 0.03
run syn150_sensitive_dpad12 "$T/syn" "$T/syn150.fq" "--sensitive --dpad 12"
run syn100_default          "$T/syn" "$T/syn100.fq" ""
run syn100_k5               "$T/syn" "$T/syn100.fq" "-k 5"
# the reference built with -DENABLE_I16 (oracle/build_ref.sh: bowtie2-align-l-i16), 16-bit DP path forced with --no-sse8
run16() { OMP_NUM_THREADS=4 "$B/bowtie2-align-l-i16" -x "$2" -U "$3" -S "$T/o.sam" -p 8 $4 2>/dev/null; grep -v '^@' "$T/o.sam" | LC_ALL=C sort > "$HERE/$1.sam"; }
run16 syn100_k5_i16   "$T/syn" "$T/syn100.fq" "-k 5 --no-sse8"
run16 syn150_sens_i16 "$T/syn" "$T/syn150.fq" "--no-sse8 --sensitive --dpad 12"
# minimum score -300: below the 8-bit range, so that build takes its 16-bit path although 8-bit SSE is enabled (aligner_sw.cpp:79)
run16 syn100_min300_i16 "$T/syn" "$T/syn100.fq" "--score-min L,-300,0 -k 3"
# reference with runs of N (record boundaries inside sequences, N at both ends of the text), 2 sequences
python "$ROOT/tools/gen_data.py" ref "$T/synN.fa" 200000 --seed 99 --repeats 0.1 --nseq 2 --ngaps 40
"$B/bowtie2-build-l" "$T/synN.fa" "$T/synN" >/dev/null 2>&1
python "$ROOT/tools/gen_data.py" reads "$T/synN.fa" "$T/synN120.fq" 1500 120 --seed 9 --randqual
run synN120_default         "$T/synN" "$T/synN120.fq" ""
# high-copy repeats (2 families x ~150 copies with 2 % divergence): seed ranges of 128+ rows, i.e. Random1toN's "seen list" mode and
# its conversion to a swap list, RowSampler masses, -D streaks, several DPs per read
python "$ROOT/tools/gen_data.py" ref "$T/synR.fa" 200000 --seed 77 --repeats 0.45 --nfam 2
"$B/bowtie2-build-l" "$T/synR.fa" "$T/synR" >/dev/null 2>&1
python "$ROOT/tools/gen_data.py" reads "$T/synR.fa" "$T/synR100.fq" 1500 100 --seed 11 --subst 0.02
run synR100_default         "$T/synR" "$T/synR100.fq" ""
run synR100_k3_vs           "$T/synR" "$T/synR100.fq" "-k 3 --very-sensitive"
# reads beyond the reference's 160-row limit: 250 / 600 / 1000 bp, through oracle/_ref/bowtie2-align-l-long = the
# -DENABLE_I16 reference with ALN_MAX_ROWS / ALN_MAX_COLS raised to 1000 / 1072 in a scratch copy (NOT the reference as shipped;
# oracle/build_ref.sh).  -p 1: that build's DP buffers are sized per read slot.  Same reads as tests/common.py reads_path("synLmix").
: > "$T/synLmix.fq"
for spec in "250 150 21 0.01" "600 120 22 0.02" "1000 120 23 0.02"; do set -- $spec
  python "$ROOT/tools/gen_data.py" reads "$T/syn.fa" "$T/synL$1.fq" $2 $1 --seed $3 --subst $4
  awk -v L=$1 'NR%4==1{sub(/^@/,"@L" L "_")}1' "$T/synL$1.fq" >> "$T/synLmix.fq"
done
runL() { OMP_NUM_THREADS=4 "$B/bowtie2-align-l-long" -x "$2" -U "$3" -S "$T/o.sam" -p 1 $4 2>/dev/null; grep -v '^@' "$T/o.sam" | LC_ALL=C sort > "$HERE/$1.sam"; }
runL synL_mix_i16     "$T/syn" "$T/synLmix.fq" ""
runL synL_mix_k3_sens "$T/syn" "$T/synLmix.fq" "--sensitive -k 3"
# 170..255 bp: the lengths at which that binary is a coherent oracle (its seed layer keeps read length and seed offsets in 8 bits)
: > "$T/synL255.fq"
for spec in "170 400 31 0.02" "200 400 32 0.02" "230 400 33 0.03" "255 400 34 0.02"; do set -- $spec
  python "$ROOT/tools/gen_data.py" reads "$T/syn.fa" "$T/synL$1.fq" $2 $1 --seed $3 --subst $4
  awk -v L=$1 'NR%4==1{sub(/^@/,"@L" L "_")}1' "$T/synL$1.fq" >> "$T/synL255.fq"
done
runL synL255_i16     "$T/syn" "$T/synL255.fq" ""
runL synL255_k3_sens "$T/syn" "$T/synL255.fq" "--sensitive -k 3"
rm -rf "$T"
ls -la "$HERE"
