# Thresholds of the E-step tiers taken from the visit count (ONI_LDA_RETIER=1, default) against the
# pair-count rule alone (=0), on what one GPU holds of a flow day cut 8 ways and of a DNS day cut 4 ways.
for spec in "flow_day 125000" "dns_day 500000" "flow_day 1000000"; do set -- $spec; for R in 0 1; do
ONI_LDA_RETIER=$R python bench.py --shape $1 --docs $2 --steps 5 --warmup 3 --no-cpu-baseline --no-cli --no-scoring 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); p=d['per_rank']
print('$1 $2 docs retier=$R: step %.2f  E %.2f  excl_end %.2f (%d docs)  long_end %.2f (%d docs)  main_end %.2f  longest %d' % (d['ms_per_step'], p['estep_ms'][0], p['end_exclusive_ms'][0], p['docs_exclusive'][0], p['end_long_ms'][0], p['docs_long'][0], p['end_main_ms'][0], p['longest_doc_pairs'][0]))"
done; done
