#!/bin/bash
# Builds tests/dropin/_ref/dropin_ref from the reference's UNMODIFIED sources where they lie (/root/reference/src) against
# the forwarding openfhe.h of sorting-fhe_b200/host/compat and libckks_b200.so.  Output only into tests/dropin/_ref/
# (git-ignored; it travels to the GPU box with the snapshot like the other built files).
set -e
cd "$(dirname "$0")"
REF=${REF:-/root/reference}
[ -d "$REF/src" ] || { echo "reference tree not present: skipping dropin_ref"; exit 0; }
ROOT=../..
mkdir -p _ref
SRCS="$REF/src/sign.cpp $REF/src/comparison.cpp $REF/src/encryption.cpp $REF/src/mehp24/mehp24_utils.cpp $REF/src/mehp24/mehp24_sort.cpp \
      $REF/src/k-way/EvalUtils.cpp $REF/src/k-way/Masking.cpp $REF/src/k-way/SortUtils.cpp $REF/src/k-way/Sorter.cpp"
g++ -O1 -std=c++17 -fopenmp -w -I $ROOT/sorting-fhe_b200/host/compat -I $REF/src -I $REF/src/k-way -I $REF/src/mehp24 \
    dropin_ref.cpp $SRCS -o _ref/dropin_ref -L $ROOT/sorting-fhe_b200 -lckks_b200 -Wl,-rpath,'$ORIGIN/../../../sorting-fhe_b200'
echo "built tests/dropin/_ref/dropin_ref"
