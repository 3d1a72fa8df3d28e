"""oni_ml_b200 -- Python host mirror of the B200-native LDA engine and event scorer
that replaces the oni-lda-c step of rabarona/oni-ml.

The product is the C-ABI library ``liboni_lda.so`` (include/oni_lda.h) and the
``lda`` executable built from ``csrc/``; the modules here bind it with ctypes for
tests and the bench and hold the host-side logic that has no arithmetic
(corpus containers, file formats, document sharding).
"""
from ._lib import LDA_BIN, LIB_PATH, OniError, Settings  # noqa: F401
