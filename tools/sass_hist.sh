#!/bin/bash
# Opcode histogram of one kernel of an object file (cuobjdump -sass): the evidence for which
# hardware paths a kernel uses (UBLKCP = bulk-copy engine, SYNCS = mbarrier, HMMA = mma.sync,
# LDSM = ldmatrix, MOVM = movmatrix, LDGSTS = cp.async; UTC*MMA / LDTM / UTMALDG would be
# tcgen05 / TMEM / tensor-map TMA).
#   tools/sass_hist.sh <object> <substring of the demangled kernel name>
set -eu
OBJ=$1; PAT=$2
cuobjdump -sass "$OBJ" | c++filt | awk -v pat="$PAT" '
  /Function :/ { on = index($0, pat) > 0; if (on) print "# " $0 }
  on && /^ +\/\*[0-9a-f]+\*\// { op = ($2 ~ /^@/) ? $3 : $2; sub(/;$/, "", op); split(op, a, "."); n[a[1]]++; total++ }
  END { for (k in n) printf "%6d %s\n", n[k], k | "sort -rn"; close("sort -rn"); printf "%6d total\n", total }'
