// E-step instantiations for G = 8 lanes per document (see estep_inst.inc).
#define ONI_G 8
#include "estep_inst.inc"
