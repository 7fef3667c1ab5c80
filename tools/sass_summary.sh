#!/bin/bash
# Opcode histogram per kernel of the built product library: the SASS mnemonics that prove tcgen05 / TMEM / TMA / bulk copies
# are on the path (B200_PROFILING.md), plus registers and instruction counts. usage: tools/sass_summary.sh > profiles/rNN_sass_summary.txt
lib=${1:-openpose_b200/lib/libopb200.so}
echo "# cuobjdump -sass $lib ($(date -u +%F)), nvcc $(nvcc --version | grep release | sed 's/.*release //')"
echo "# kernel | instructions | UTCHMMA (tcgen05.mma) [.2CTA] | LDTM (tcgen05.ld) | UTMALDG (TMA load) | UTMASTG (TMA store) | UBLKCP (bulk copy) | UTCBAR (tcgen05.commit) | SYNCS (mbarrier) | HMMA (mma.sync) | R2UR"
cuobjdump -sass "$lib" | awk '
  /Function : /{ if (name != "") out(); name=$3; n=0; split("",c) }
  /^[[:space:]]+\/\*[0-9a-f]+\*\//{ n++; op=$2; if (op ~ /^@/) op=$3; sub(/;$/,"",op);
      if (op ~ /^UTCHMMA/) { c["mma"]++; if (op ~ /2CTA/) c["mma2"]++ }
      if (op ~ /^LDTM/) c["ldtm"]++; if (op ~ /^UTMALDG/) c["tmald"]++; if (op ~ /^UTMASTG/) c["tmast"]++;
      if (op ~ /^UBLKCP/) c["blk"]++; if (op ~ /^UTCBAR/) c["bar"]++; if (op ~ /^SYNCS/) c["syncs"]++;
      if (op ~ /^HMMA/) c["hmma"]++; if (op ~ /^R2UR/) c["r2ur"]++ }
  function out() { printf "%s | %d | %d [%d] | %d | %d | %d | %d | %d | %d | %d | %d\n", name, n, c["mma"], c["mma2"], c["ldtm"], c["tmald"], c["tmast"], c["blk"], c["bar"], c["syncs"], c["hmma"], c["r2ur"] }
  END{ if (name != "") out() }' | c++filt | sed 's/(opb::ConvParams)//; s/void //'
