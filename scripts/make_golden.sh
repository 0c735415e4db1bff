#!/usr/bin/env bash
# make_golden.sh - pins the oracle to the REAL reference (Seiichi-Yahiro/point-cloud, point-converter crate).
#
# Cannot run in the build image (no cargo / rustc, no network); run it on any machine that has a Rust toolchain
# and a checkout of the reference:
#
#     REFERENCE=/path/to/point-cloud scripts/make_golden.sh
#
# What it does:
#   1. dumps two synthetic clouds as binary little-endian PLY (float x, y, z; uchar red, green, blue) with
#      scripts/golden_tool.py (numpy only: the generator is integer arithmetic up to the final int -> float
#      conversion, so every machine produces the same bytes; the PLY's own SHA-256 is recorded):
#        c1_cube_1m      1 000 000 uniform points in the unit cube          (BASELINE configs[0])
#        multiroot_200k  200 000 points over 3 x 2 x 1 level-0 cells incl. negative indices, 2.5-D
#   2. builds the reference's own CLI (cargo build --release -p point-converter) and converts each PLY into a
#      fresh directory with default settings (batch 10 000, config 5000 / 96 / 1000.0),
#   3. writes tests/golden/<name>.json: for every node file the SHA-256 of its CANONICAL form (48-byte header,
#      grid records sorted bytewise, overflow entries sorted by child index with their lists in file order - the
#      orders the reference leaves to HashMap iteration) plus metadata.json verbatim.
# tests/test_golden.py then checks the ORACLE (and, on a GPU box, the CUDA path) against those files; without them
# it reports "golden absent" (parity stays unpinned).
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REFERENCE="${REFERENCE:?set REFERENCE to a checkout of Seiichi-Yahiro/point-cloud}"
WORK="${WORK:-$(mktemp -d)}"
command -v cargo >/dev/null || { echo "cargo not found: run this on a machine with a Rust toolchain" >&2; exit 2; }
(cd "$REFERENCE" && cargo build --release -p point-converter)
BIN="$REFERENCE/target/release/point-converter"
[ -x "$BIN" ] || BIN="$REFERENCE/target/release/point_converter"
for name in c1_cube_1m multiroot_200k; do
    python "$ROOT/scripts/golden_tool.py" ply "$name" "$WORK/$name.ply"
    rm -rf "$WORK/$name.out"
    "$BIN" -o "$WORK/$name.out" -f "$WORK/$name.ply"
    python "$ROOT/scripts/golden_tool.py" digest "$name" "$WORK/$name.ply" "$WORK/$name.out" "$ROOT/tests/golden/$name.json"
    echo "wrote tests/golden/$name.json"
done
