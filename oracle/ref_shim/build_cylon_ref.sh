#!/usr/bin/env bash
# Compiles the subset of the reference that builds unmodified against pyarrow's Arrow C++ (24.x) into
# oracle/_ref/libcylon_ref.so: hash partition kernels, row comparators, hash and sort joins, murmur3, and the table
# operators (table.cpp, groupby/, partition/, the Arrow all-to-all) driven through an in-process communicator. Sources are compiled
# IN PLACE from /root/reference (never copied). The reference's own build system (cmake + Arrow 2.0 + MPI + glog
# fetched from the network) cannot run here; see DESIGN.md for what does not compile and why.
set -euo pipefail
OUT="$1"
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=/root/reference/cpp/src/cylon
PA_INC=$(python -c "import pyarrow as pa; print(pa.get_include())")
PA_LIB=$(python -c "import pyarrow as pa; print(pa.get_library_dirs()[0])")
LIBARROW=$(ls "$PA_LIB"/libarrow.so.* | head -1)
LIBARROW_COMPUTE=$(ls "$PA_LIB"/libarrow_compute.so.* | head -1)
SRCS="$REF/join/join.cpp $REF/join/hash_join.cpp $REF/join/sort_join.cpp $REF/join/join_utils.cpp \
      $REF/arrow/arrow_partition_kernels.cpp $REF/arrow/arrow_comparator.cpp $REF/util/arrow_utils.cpp $REF/arrow/arrow_kernels.cpp \
      $REF/util/murmur3.cpp $REF/util/copy_arrray.cpp \
      $REF/table.cpp $REF/table_api.cpp $REF/column.cpp $REF/row.cpp $REF/indexing/index.cpp $REF/partition/partition.cpp \
      $REF/groupby/groupby.cpp $REF/groupby/hash_groupby.cpp $REF/arrow/arrow_all_to_all.cpp $REF/net/ops/all_to_all.cpp \
      $REF/net/TxRequest.cpp $REF/util/builtins.cpp $REF/io/csv_read_config.cpp $REF/io/csv_write_config.cpp"
if [ -f "$OUT/libcylon_ref.so" ] && [ "$OUT/libcylon_ref.so" -nt "$HERE/cylon_ref_shim.cpp" ] && \
   [ "$OUT/libcylon_ref.so" -nt "$HERE/stubs/compat.h" ] && [ "$OUT/libcylon_ref.so" -nt "$0" ]; then exit 0; fi
# one object per source, in parallel (objects live under build/, outside the tree that travels), then one link
OBJ="$HERE/../../build/ref_obj"
mkdir -p "$OBJ"
FLAGS="-std=c++20 -O2 -fPIC -pthread -w -include $HERE/stubs/compat.h -I$HERE/stubs -I$REF -I$PA_INC"
compile_one() {
  src="$1"
  obj="$OBJ/$(echo "${src#/}" | tr '/' '_' | sed 's/\.cpp$/.o/')"
  g++ $FLAGS -c "$src" -o "$obj"
}
export -f compile_one
export OBJ FLAGS
printf '%s\n' "$HERE/cylon_ref_shim.cpp" $SRCS | xargs -P "$(nproc)" -I{} bash -c 'compile_one "$@"' _ {}
g++ -shared -pthread -o "$OUT/libcylon_ref.so" "$OBJ"/*.o "$LIBARROW_COMPUTE" "$LIBARROW" -Wl,-rpath,"$PA_LIB"
echo "build_cylon_ref: built $OUT/libcylon_ref.so"
