#!/bin/bash
# tools/gpu_round_ab.sh: one-call A/B of the compile-time variants of the fan kernel (run under gpurun after building them here):
#   for v in "ta -DB2F_FAN_TILE_AHEAD=888" "pf2 -DB2F_FAN_PREFETCH=2" "vote -DB2F_FAN_UNIFORM_PIVOT=1" "n2 -DB2F_RCP_CUBIC=0" \
#            "wa -DB2F_FAN_WORD_AHEAD=1" "r64 -DB2F_FAN_ROWS=64"; do set -- $v; n=$1; shift; tools/build_variant.sh $n "$@"; done
#   gpurun --timeout 400 -- 'bash tools/gpu_round_ab.sh base ta pf2 vote n2 wa r64'
# then the parity suite on the build that is to become the default (B2F_LIBRARY selects it).  Results: profiles/rNN_fan_variants.txt.
bash tools/gpu_variants.sh "$@"
