#!/bin/bash
# SASS instruction count per kernel of libb200pt.so (code size: the traversal kernels must stay I-cache friendly)
cuobjdump -sass "${1:-pathtrace_b200/libb200pt.so}" 2>/dev/null | awk '
/Function :/ { if (name != "") print n, name; name = $3; n = 0 }
/^ +\/\*[0-9a-f]+\*\/ / { n++ }
END { print n, name }' | sed -E 's/_ZN2pt[0-9]*_GLOBAL__N__[a-z0-9_]*[0-9]+//' | sort -rn
