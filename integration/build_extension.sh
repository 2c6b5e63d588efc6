#!/bin/bash
# Builds, out of tree, (1) stock DuckDB + the reference's own rpt extension and (2) the same extension with the GPU
# body, and collects what the end-to-end tests need under baseline/_ref/bin/ (git-ignored; travels to the GPU box):
#   duckdb                        the DuckDB shell (v1.4.4, the reference's submodule)
#   stock/rpt.duckdb_extension    the unmodified reference extension (CPU)
#   gpu/rpt.duckdb_extension      integration/bloom_filter_gpu.cpp + every other reference source unmodified
# (DuckDB derives the entry point from the file name, so both files are called rpt.duckdb_extension.)
# Needs /root/reference (read-only) and ~10 minutes the first time; the second configuration re-uses the DuckDB
# objects and only re-compiles the extension.
set -euo pipefail
REPO=$(cd "$(dirname "$0")/.." && pwd)
REF=${RPT_REFERENCE_DIR:-/root/reference}
B=$REPO/baseline/_ref/build
mkdir -p "$B" "$REPO/baseline/_ref/bin/stock" "$REPO/baseline/_ref/bin/gpu"
common=(-G Ninja -DEXTENSION_STATIC_BUILD=1 -DBUILD_EXTENSIONS=tpch -DOVERRIDE_GIT_DESCRIBE=v1.4.4-0-g8827746300
        -DBUILD_UNITTESTS=0 -DENABLE_UNITTEST_CPP_TESTS=FALSE -DCMAKE_BUILD_TYPE=Release -S "$REF/duckdb" -B "$B")
if [ ! -f "$REPO/baseline/_ref/bin/stock/rpt.duckdb_extension" ]; then
	cmake "${common[@]}" -DDUCKDB_EXTENSION_CONFIGS="$REF/extension_config.cmake"
	cmake --build "$B" -j "${JOBS:-6}"
	cp "$B/duckdb" "$REPO/baseline/_ref/bin/duckdb"
	cp "$B/extension/rpt/rpt.duckdb_extension" "$REPO/baseline/_ref/bin/stock/rpt.duckdb_extension"
fi
python -c "import sys; sys.path.insert(0, '$REPO'); import rpt_b200; rpt_b200.build()"
cmake "${common[@]}" -DDUCKDB_EXTENSION_CONFIGS="$REPO/integration/extension_config.cmake" -DRPT_REFERENCE_DIR="$REF"
cmake --build "$B" -j "${JOBS:-6}" --target rpt_loadable_extension
cp "$B/extension/rpt/rpt.duckdb_extension" "$REPO/baseline/_ref/bin/gpu/rpt.duckdb_extension"
ls -la "$REPO/baseline/_ref/bin" "$REPO/baseline/_ref/bin/gpu"
