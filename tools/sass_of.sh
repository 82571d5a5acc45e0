#!/bin/bash
# Line-annotated SASS of one kernel of the built library (no GPU needed): tools/sass_of.sh <mangled-name substring> [build dir]
# e.g. tools/sass_of.sh k_neeILb0ELb1ELb1ELb0ELb0 > /tmp/nee.sass     (then grep "pt_intersect.h\", line 5" ...)
PAT=$1; BUILD=${2:-$(cd $(dirname $0)/.. && pwd)/pathtrace_b200/csrc/build}
TMP=$(mktemp -d); trap "rm -rf $TMP" EXIT
(cd $TMP && cuobjdump -xelf all $(realpath $BUILD)/kernels.o > /dev/null)
nvdisasm -g $TMP/*.cubin | awk -v pat="$PAT" '
  /^\.text\./ { on = index($0, pat) > 0 }
  /^\.section/ { on = 0 }
  on { print }' | grep -v "inlined at" | sed 's#//\#\# File ".*/\([^/"]*\)", line \([0-9]*\)#// \1:\2#'
