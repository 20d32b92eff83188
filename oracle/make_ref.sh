#!/bin/sh
# Stage the UNMODIFIED reference for the GPU box (test infrastructure; no reference file enters git).
#
# /root/reference exists only in the build container.  The reference is Python, so "building" it
# for the box means packing the files the drop-in tests and the reference arm of bench.py run --
# mdistiller/ (package), configs/ (yaml), tools/train.py (entry script) -- into ONE archive under the
# git-ignored oracle/_ref/, which travels to the box with the built .so.  mdistiller-ddp_b200/refhost.py
# unpacks it into a temporary directory at run time.  A manifest of SHA-256 sums is written beside it
# so a test can show that what ran on the box is byte-for-byte what is mounted here.
set -eu
REF="${MDISTILLER_REF_ROOT:-/root/reference}"
OUT="$(cd "$(dirname "$0")" && pwd)/_ref"
[ -f "$REF/mdistiller/distillers/DKD.py" ] || { echo "no reference tree at $REF" >&2; exit 1; }
mkdir -p "$OUT"
(cd "$REF" && find mdistiller configs tools/train.py LICENSE -type f ! -name '*.pyc' | LC_ALL=C sort) > "$OUT/files.txt"
(cd "$REF" && xargs sha256sum < "$OUT/files.txt") > "$OUT/MANIFEST.sha256"
tar -C "$REF" -czf "$OUT/mdistiller_ref.tar.gz" -T "$OUT/files.txt"
rm -f "$OUT/files.txt"
echo "staged $(wc -l < "$OUT/MANIFEST.sha256") reference files -> $OUT/mdistiller_ref.tar.gz"
