# A/B of the intersection-kernel variants (development aid, GPU box): bash tools/ab_variants.sh "<variants>" "<scenes>" tag
set -x
VARS=${1:-"-1 16 216"}; SCENES=${2:-"2 3 4"}; TAG=${3:-ab}
for v in $VARS; do for s in $SCENES; do
GPRT_ISECT_VARIANT=$v python tools/profile_frame.py --scene $s --no-oracle > gpurun_out/${TAG}_v${v}_s$s.json 2>&1
GPRT_ISECT_VARIANT=$v python tools/profile_frame.py --scene $s --no-oracle --fast > gpurun_out/${TAG}_v${v}_s${s}_fast.json 2>&1
done; done
