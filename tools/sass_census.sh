#!/usr/bin/env bash
# SASS mnemonic census of the shipped objects (vllm_splitwise_b200/build/*.o) -> profiles/rNN_sass_census.txt
#   bash tools/sass_census.sh > profiles/r05_sass_census.txt
cd "$(dirname "$0")/.."
echo "# cuobjdump -sass census of vllm_splitwise_b200/build/*.o (sm_100a), $(git rev-parse --short HEAD), $(date -u +%F)"
echo "# UBLKCP = cp.async.bulk (TMA, non-tensor), UTMALDG = tensor-map TMA load, UTCHMMA = tcgen05.mma (f16 kind), LDTM/STTM = tcgen05.ld/st,"
echo "# UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, HMMA = mma.sync, LDSM = ldmatrix, LDGSTS = cp.async (16 B), ACQBULK/...: bulk-group waits"
printf "%-34s %7s %7s %7s %6s %6s %6s %6s %6s %6s %7s\n" object UBLKCP UTMALDG UTCHMMA LDTM STTM UTCBAR SYNCS HMMA LDSM LDGSTS
for o in vllm_splitwise_b200/build/*.o; do
  t=$(mktemp); cuobjdump -sass "$o" > "$t" 2>/dev/null
  c() { grep -c "$1" "$t"; }
  printf "%-34s %7s %7s %7s %6s %6s %6s %6s %6s %6s %7s\n" "$(basename $o)" "$(c UBLKCP)" "$(c UTMALDG)" "$(c UTCHMMA)" "$(c 'LDTM')" "$(c 'STTM')" "$(c UTCBAR)" "$(c SYNCS)" "$(c 'HMMA')" "$(c 'LDSM')" "$(c LDGSTS)"
  rm -f "$t"
done
