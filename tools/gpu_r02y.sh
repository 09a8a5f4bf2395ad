#!/bin/bash
# round 2: ncu --set full of the side kernels on the final tree (one launch each, application replay where the kernel
# works in place)
set -x
mkdir -p gpurun_out
timeout 300 python tools/ncu_extras.py config3 stable locate next > gpurun_out/r02y_extras_plain.log 2>&1; echo "plain exit $?"; tail -5 gpurun_out/r02y_extras_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"^match_kernel|interaction_offsets_kernel" -c 2 -o gpurun_out/r02y_locate python tools/ncu_extras.py locate > gpurun_out/r02y_ncu_locate.log 2>&1; echo "ncu locate exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"unknown_step_kernel" -c 1 -o gpurun_out/r02y_unknown python tools/ncu_extras.py config3 > gpurun_out/r02y_ncu_unknown.log 2>&1; echo "ncu unknown exit $?"
timeout 900 ncu --set full --clock-control none --replay-mode application -k regex:"stable_pass_kernel" -c 1 -o gpurun_out/r02y_stable python tools/ncu_extras.py stable > gpurun_out/r02y_ncu_stable.log 2>&1; echo "ncu stable exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"hash_kernel|^interaction_kernel|weld_interaction" -c 3 -o gpurun_out/r02y_next python tools/ncu_extras.py next > gpurun_out/r02y_ncu_next.log 2>&1; echo "ncu next exit $?"
ls -la gpurun_out/r02y*
