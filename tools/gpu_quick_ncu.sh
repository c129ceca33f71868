#!/bin/bash
# tools/gpu_quick_ncu.sh -- MPC parity tests + bench line, then one full ncu capture of the MPC kernel (one gpurun call).
set -u
TAG=${1:-q}
bash tools/gpu_quick.sh $TAG
bash tools/gpu_ncu_mpc.sh ${TAG}_ncu ${2:-mpc_v3_kernel}
