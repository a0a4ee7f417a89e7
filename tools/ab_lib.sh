# same-box A/B of two builds of the library: tools/ab_lib.sh <other .so>   (bench legs without training / cpu baseline)
OTHER=$1
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],4), round(r["us_per_launch"],2), round(r["us_per_launch_l2_flushed_single"],2), [round(x["ms_per_step"],3) for x in d["c5_sweep"]["rows"]], round(d["varying_geometry"]["stream"]["ms_per_step"],4))'
for i in 1 2; do
echo -n "tree  "; timeout 300 python bench.py --no-train --no-cpu-baseline 2>/dev/null | python -c "$P"
echo -n "other "; SKG_LIB_PATH=$OTHER timeout 300 python bench.py --no-train --no-cpu-baseline 2>/dev/null | python -c "$P"
done
