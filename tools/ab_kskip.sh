for rep in 1 2; do
for v in k0 k1 cur; do
  if [ $v = cur ]; then unset UZ_LIB_PATH; else export UZ_LIB_PATH=build/libuz_$v.so; fi
  echo "== $v: $(python tools/bench_conv.py 2>&1 | grep -E 'stem|c3 64->64|c3 128->64|dg 64->128|c3 64->128|total' | awk '{printf "%s %s=%s  ", $1, $2, $(NF-6)}')"
done
done
for v in k0 k1 cur k0 k1 cur; do
  if [ $v = cur ]; then unset UZ_LIB_PATH; else export UZ_LIB_PATH=build/libuz_$v.so; fi
  python bench.py --no-secondary --no-cpu-baseline 2>/dev/null | python -c "import json,sys; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v unet', round(l['value'],1), round(l['ms_per_step'],3), l['clocks']['sm_mhz'])"
done
