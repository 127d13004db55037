#!/bin/bash
# Every experiment of EXPERIMENTS.md is a build flag: this compiles the library once per flag group (no GPU needed) so that a
# rejected experiment cannot rot.  ~1 minute per group.
set -u
fail=0
for f in "-DLZB_X_EARLY_PATH" "-DLZB_X_CARRY" "-DLZB_X_PBC2 -DLZB_X_CVAL -DLZB_X_SPARSE_SUM" "-DLZB_X_BRANCHFREE_DIV -DLZB_TABLES_IN_CONSTANT" \
         "-DLZB_X_NO_GATHER8 -DLZB_X_NO_GMZ_STATS -DLZB_X_NO_SHARED_TABLES -DLZB_X_SAMPLER_ONE_PER_PAIR" "-DLZB_X_GMZ_SPARSE" \
         "-DLZB_X_ROW_PREFETCH=1" "-DLZB_X_ROW_PREFETCH=2 -DLZB_X_ROW_PREFETCH_AT_DESCENT" "-DLZB_X_GATHER_PREFETCH -DLZB_X_GATHER_STCS -DLZB_X_GATHER_LDG" \
         "-DLZB_X_GATHER_LDNC" "-DLZB_X_GATHER_LDCG" "-DLZB_X_GATHER_LDLU" "-DLZB_X_SCALAR_PAIR -DLZB_X_SCALAR_BALANCED" "-DLZB_X_SCALAR_PAIR -DLZB_SCALAR_PAIR_MIN_BLOCKS=4" \
         "-DLZB_X_SCALAR_WARP_TICKET -DLZB_X_SCALAR_PRECISE_DIV" "-DLZB_SCALAR_THREADS=128" "-DLZB_X_SCALAR_PIPELINED" "-DLZB_X_SCALAR_ROWLOADS" \
         "-DLZB_X_NO_LONE_WALKER" "-DLZB_X_LONE_WALKER_ALL -DLZB_X_LONE_PREFETCH" "-DLZB_X_NO_REDUX -DLZB_X_NO_LONE_STAGE" "-DLZB_X_LONE_STAGE_LSU" \
         "-DLZB_X_PAIR_SPLIT -DLZB_X_LONE_EARLY_RELEASE" "-DLZB_CHECKED" "-DLZB_TIMELINE" "-DLZB_MIN_BLOCKS=6"; do
  if python scripts/build_variant.py flagcheck $f > /tmp/lzb_flagcheck.log 2>&1; then echo "OK   $f"; else echo "FAIL $f"; grep -i " error" /tmp/lzb_flagcheck.log | head -n 3; fail=1; fi
done
rm -f lightzero_b200/liblzb200_flagcheck.so
exit $fail
