#!/bin/bash
# A/B timing of the working tree (B) against HEAD (A) on ONE GPU box, alternating runs:
#   tools/ab_compare.sh [populations...]          (default: 8192 65536)
# Box-to-box differences are a few percent; within one box the kernel times repeat to ~0.1 %.
set -e
cd "$(dirname "$0")/.."
POPS="${@:-8192 65536}"
if git diff --quiet; then echo "working tree equals HEAD: nothing to compare"; exit 1; fi
git stash -q
python __graft_entry__.py build >/dev/null 2>&1 && cp gaudi_b200/libgaudimm_b200.so gaudi_b200/libA.bin
git stash pop -q
python __graft_entry__.py build >/dev/null 2>&1 && cp gaudi_b200/libgaudimm_b200.so gaudi_b200/libB.bin
/usr/local/graft/bin/gpurun --timeout 500 -- "for r in 1 2; do for v in A B; do cp gaudi_b200/lib\$v.bin gaudi_b200/libgaudimm_b200.so; echo -n \"\$v \"; timeout 100 python tools/latency_small_pop.py $POPS 2>&1 | cut -c1-150; done; done; cp gaudi_b200/libB.bin gaudi_b200/libgaudimm_b200.so; timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2" 2>&1 | tail -14
rm -f gaudi_b200/libA.bin gaudi_b200/libB.bin
