#!/bin/bash
# Opcode counts per kernel family of the built library (cuobjdump -sass): evidence for the tcgen05 / TMEM / TMA / dp4a claims.
# usage: tools/sass_summary.sh > profiles/<round>_sass_summary.txt
cd "$(dirname "$0")/../llama-mi50.cpp_b200/csrc/build"
count() { grep -c -E "$2" "$1" || true; }
for f in k_gemv_stream_i_q4k_q6k k_mmq_tc k_fattn k_mmvq k_quantize; do
  cuobjdump -sass $f.o > /tmp/sass_$f.txt 2>/dev/null
  echo "== $f.o  ($(grep -c 'Function :' /tmp/sass_$f.txt) kernels, $(grep -c -E '^\s+/\*[0-9a-f]{4,5}\*/' /tmp/sass_$f.txt) instructions)"
  printf "   UTCHMMA (tcgen05.mma) %s  UTCBAR/UTCCOMMIT %s  LDTM (tcgen05.ld) %s  UTMALDG (TMA tensor) %s  UBLKCP (TMA bulk) %s  SYNCS (mbarrier) %s\n" \
    $(count /tmp/sass_$f.txt UTCHMMA) $(count /tmp/sass_$f.txt 'UTCBAR|UTCCOMMIT') $(count /tmp/sass_$f.txt LDTM) $(count /tmp/sass_$f.txt UTMALDG) $(count /tmp/sass_$f.txt UBLKCP) $(count /tmp/sass_$f.txt SYNCS)
  printf "   IDP.4A (dp4a) %s  HMMA (mma.sync f16) %s  IMMA %s  LDSM (ldmatrix) %s  LDGSTS (cp.async) %s  USETMAXREG %s  SHFL %s  BAR %s\n" \
    $(count /tmp/sass_$f.txt 'IDP.4A') $(count /tmp/sass_$f.txt HMMA) $(count /tmp/sass_$f.txt IMMA) $(count /tmp/sass_$f.txt LDSM) $(count /tmp/sass_$f.txt LDGSTS) $(count /tmp/sass_$f.txt USETMAXREG) $(count /tmp/sass_$f.txt SHFL) $(count /tmp/sass_$f.txt 'BAR\.')
done
