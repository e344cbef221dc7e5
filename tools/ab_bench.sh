# usage: ab.sh <alt lib> <label>   -- two alternating runs of the main bench step with each library
for i in 1 2; do
  for L in "" "$1"; do
    LSB200_LIB=$L python bench.py --steps 20 --no-cpu-baseline --no-xt32 --no-other-configs 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$L' or 'default', round(d['value']), round(d['ms_per_step'], 4), d['stages_ms'], 'parity', d.get('parity_check', {}).get('mismatch_tri'), d.get('parity_check', {}).get('mismatch_t'))"
  done
done
