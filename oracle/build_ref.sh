#!/usr/bin/env bash
# Builds oracle/_ref/libbert_ref.so = the reference's own CPU implementation of
# the bert_eval path, compiled from the sources where they lie under
# /root/reference (never copied into this repo).  Test infrastructure only.
#
#   sources : bert.cpp + ggml/src/{ggml,ggml-alloc,ggml-backend,ggml-quants}.c
#   flags   : those of the reference's CMakeLists.txt:118-119,164-177
#             (-O3 -mavx -mavx2 -mfma -mf16c) and ggml/CMakeLists.txt:125-127
#   shim    : oracle/ref_shim/ instead of tokenizer.cpp + tokenizers-cpp (Rust)
#   patches : three one-token edits applied by sed to a TEMPORARY copy of
#             bert.cpp (deleted afterwards); none changes arithmetic:
#       P1  bert.cpp:1085  nodes[n_nodes-2] -> nodes[n_nodes-1]
#           (bert_eval_batch copies the scalar 1/|e| node instead of the
#            embedding; n_nodes-1 is the embedding, i.e. `res` at :1084)
#       P2a bert.cpp:826   1 GiB compute-arena cap -> 64 GiB
#       P2b bert.cpp:814   arena head-room 1.1 -> 4.0 (O(N^2) attention
#           tensors are not in the 4-token dry run; N>=384 aborts otherwise)
#
#   VARIANT=portable ./build_ref.sh  builds oracle/_ref/libbert_ref_portable.so instead: the same sources
#   WITHOUT -mavx -mavx2 -mfma -mf16c, i.e. the reference's own generic-C code paths (the `#else` branches
#   of ggml-quants.c:632-700,2426-2523 / ggml.c that it runs on a CPU without AVX2).  Used only to
#   measure how far two builds of the REFERENCE ITSELF are apart (tests/golden: emb_portable): the Q8
#   activation rounding makes deep Q4 models chaotic, so that spread is the floor of any re-implementation.
set -euo pipefail
REF=${REF:-/root/reference}
VARIANT=${VARIANT:-avx2}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
if [ ! -f "$REF/bert.cpp" ]; then
  echo "build_ref: $REF not present; keeping prebuilt $OUT (if any)"; exit 0
fi
mkdir -p "$OUT"
TMP=$(mktemp -d)
trap 'rm -rf "$TMP"' EXIT
SIMD="-mavx -mavx2 -mfma -mf16c"; LIBNAME=libbert_ref.so
if [ "$VARIANT" = "portable" ]; then SIMD=""; LIBNAME=libbert_ref_portable.so; fi
CF="-O3 $SIMD -D_GNU_SOURCE -D_XOPEN_SOURCE=600 -fPIC -DNDEBUG \
    -I$REF/ggml/include -I$REF/ggml/include/ggml -I$REF/ggml/src"
pids=()
for f in ggml ggml-alloc ggml-backend ggml-quants; do
  gcc -std=c11 $CF -w -c "$REF/ggml/src/$f.c" -o "$TMP/$f.o" & pids+=($!)
done
sed -e 's/gf->nodes\[gf->n_nodes - 2\]/gf->nodes[gf->n_nodes - 1]/' \
    -e 's/int64_t GB = 1 << 30;/int64_t GB = (int64_t)64 << 30;/' \
    -e 's/1\.1 \* (new_bert->mem_per_token \* N)/4.0 * (new_bert->mem_per_token * N)/' \
    "$REF/bert.cpp" > "$TMP/bert_patched.cpp"
ndiff=$(diff "$REF/bert.cpp" "$TMP/bert_patched.cpp" | grep -c '^>' || true)
[ "$ndiff" = "3" ] || { echo "build_ref: expected 3 patched lines, got $ndiff"; exit 1; }
g++ -std=c++17 $CF -w -include unordered_map -include mutex -include array -include stdexcept \
    -I"$REF" -I"$HERE/ref_shim" -c "$TMP/bert_patched.cpp" -o "$TMP/bert.o" & pids+=($!)
g++ -std=c++17 $CF -w -I"$REF" -I"$HERE/ref_shim" -c "$HERE/ref_shim/tok_shim.cpp" -o "$TMP/tok_shim.o" & pids+=($!)
for p in "${pids[@]}"; do wait "$p"; done
g++ -shared -o "$OUT/$LIBNAME" "$TMP"/*.o -lpthread -lm
echo "build_ref: wrote $OUT/$LIBNAME"
