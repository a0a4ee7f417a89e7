# same-box A/B of library builds on the fused fc_2 launch: tools/ab_gemm.sh <lib.so> [<lib.so> ...]
for i in 1 2; do
echo "tree"; timeout 120 python tools/gemm_bench.py --fused-only 2>&1 | grep -E "^fused4|^plain" | tail -4
for L in "$@"; do echo "$L"; SKG_LIB_PATH=$L timeout 120 python tools/gemm_bench.py --fused-only 2>&1 | grep -E "^fused4|^plain" | tail -4; done
done
