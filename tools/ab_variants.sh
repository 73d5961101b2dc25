for v in 0 1 8 16 24; do
  echo "variant $v"; GPRT_ISECT_VARIANT=$v python tools/profile_frame.py --scene 2 --no-oracle --reps 6 | python -c "import json,sys; d=json.load(sys.stdin); print('parity', d['plain_dispatch_ms'])"
  GPRT_ISECT_VARIANT=$v python tools/profile_frame.py --scene 2 --no-oracle --reps 6 --fast | python -c "import json,sys; d=json.load(sys.stdin); print('fast  ', d['plain_dispatch_ms'])"
done
