#!/bin/bash
# Static opcode histogram of one kernel in the built library.  usage: tools/sass_count.sh [lib] [kernel-substring] [top]
lib=${1:-raytracer-challenge_b200/csrc/librtc_b200.so}; pat=${2:-render_f32_small_kernelILi1ELb0}
cuobjdump -sass "$lib" | awk -v pat="$pat" '/Function :/ {on = index($0, pat) > 0} on && /^[ \t]+\/\*[0-9a-f]+\*\//' \
  | sed -E 's/^[ \t]+\/\*[0-9a-f]+\*\/[ \t]+(@!?U?P[0-9T]+[ \t]+)?([A-Z0-9_]+).*/\2/' | sort | uniq -c | sort -rn | awk '{t+=$1; print} END {print t, "total"}' | head -${3:-16}
