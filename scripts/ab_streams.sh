# A/B of the host-side stream branches (decoder.py): priority of the branch streams, decoder-layer stream on/off
for cfg in "0 0" "-1 0" "0 1" "-1 1" "0 1"; do
  set -- $cfg
  BIST_BRANCH_PRIO=$1 BIST_DECODER_STREAM=$2 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; j=json.loads(sys.stdin.read()); print('prio=$1 dec=$2', round(j['ms_per_step'],4), [(k, round(j[k].get('ms_per_step', j[k].get('ms_per_batch')),3)) for k in ('train','train_full','train_full_audio','decode')])"
done
