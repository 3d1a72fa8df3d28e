# E-step tiers on a 125k-document corpus (what one of 8 GPUs holds of a flow day): where the three launches end
# for different thresholds.  usage: bash profiles/tier_sweep.sh "<long_min values>" "<excl_min values>"
for L in ${1:-256 360 512}; do for X in ${2:-800 1138 1600}; do
ONI_LDA_LONG_DOC=$L ONI_LDA_EXCL_DOC=$X python bench.py --docs ${DOCS:-125000} --steps 5 --warmup 3 --no-cpu-baseline --no-cli --no-scoring 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); p=d['per_rank']
print('long>=$L excl>=$X: step %.2f  E %.2f  excl_end %.2f (%d docs)  long_end %.2f (%d docs)  main_end %.2f  longest %d' % (d['ms_per_step'], p['estep_ms'][0], p['end_exclusive_ms'][0], p['docs_exclusive'][0], p['end_long_ms'][0], p['docs_long'][0], p['end_main_ms'][0], p['longest_doc_pairs'][0]))"
done; done
