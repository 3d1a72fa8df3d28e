#!/bin/bash
# Self-check build of the long-document kernel's cp.async ring (compute-sanitizer is closed on this
# pool): every word visit compares what the ring delivered (pair count, log-beta row, old phi row)
# with plain loads of the same words and raises a flag on the first difference.  Run on a GPU box:
#   bash profiles/ring_check.sh > gpurun_out/ring_check.txt 2>&1
set -e
cd "$(dirname "$0")/.."
mkdir -p oni_ml_b200/variants
make -C oni_ml_b200/csrc -j BUILD=build_check OUTLIB=../variants/liboni_lda_check.so OUTBIN=../variants/lda_check \
     EXTRA=-DONI_RING_CHECK > /dev/null
export ONI_LDA_LIB=$PWD/oni_ml_b200/variants/liboni_lda_check.so
echo "== long-document tests on the self-check build ($ONI_LDA_LIB)"
python -m pytest tests/test_gpu_lda.py -q -m gpu -k "long_document or leaves_the or flow_day_size or config1_full"
echo "== a single 8000-pair document, 3 EM iterations"
python profiles/longdoc_single.py

echo "== negative control: the same build with the next visit's row read from the wrong ring slot (-DONI_RING_CHECK_BREAK) must FAIL the check"
make -C oni_ml_b200/csrc -j BUILD=build_check2 OUTLIB=../variants/liboni_lda_check2.so OUTBIN=../variants/lda_check2 \
     EXTRA="-DONI_RING_CHECK -DONI_RING_CHECK_BREAK" > /dev/null
ONI_LDA_LIB=$PWD/oni_ml_b200/variants/liboni_lda_check2.so python profiles/longdoc_single.py 2>&1 | tail -2 || true
